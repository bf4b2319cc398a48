"""Parity of the node setup steps on the device (plc_setup.cu) against the oracle's restatement of
virtual_lidar.cpp:209-340 and :550-625, through the C ABI (plc_vl_generate_cells, plc_vl_compute_normals,
plc_vl_generate_candidates).

Bars: cell positions, the kept set and its order, the lattice bounds, candidate x/y/z and the kept candidates are
bit-identical.  Cell normals: the double sum of the neighbours' normals runs in another order than the reference's
KD-tree result order, so a float32 component may differ in its last bit (tolerance 1.2e-7 absolute, count printed).
Pitch/yaw: device atan2 against libm, 4e-15 absolute.  PCA normals are not pinned upstream (PCL's eigen solver):
compared by angle."""
import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import _abi as A
from pointcloud_processor_b200 import context as ctxmod
from pointcloud_processor_b200 import scenes

pytestmark = pytest.mark.gpu

BOUND_FIELDS = ("min_x", "max_x", "min_y", "max_y", "min_z", "max_z", "width", "height")


def same_bounds(a, b):
    return all(getattr(a, f) == getattr(b, f) for f in BOUND_FIELDS)


def normals_of(cells):
    return np.stack([cells["nx"], cells["ny"], cells["nz"]], 1)


def assert_cells_match(got, ref, what=""):
    assert got.size == ref.size, (what, got.size, ref.size)
    for k in ("x", "y", "z"):
        assert np.array_equal(got[k].view(np.uint64), ref[k].view(np.uint64)), (what, k)
    d = np.abs(normals_of(got).astype(np.float64) - normals_of(ref).astype(np.float64))
    differing = int(np.count_nonzero(d))
    assert d.max(initial=0.0) <= 1.2e-7, (what, d.max())
    return differing


@pytest.mark.parametrize("shape,res,layers", [("L", 0.1, 10), ("rect", 0.1, 10), ("L", 0.2, 3), ("rect", 0.05, 24)])
def test_cell_lattice_matches_oracle(ctx, shape, res, layers):
    area, normals = scenes.excavation_area(shape, 0.05)
    got, gb = ctx.generate_cells(area, normals, res, layers, 1.5)
    ref, rb = O.generate_cells(area, normals, res, layers, 1.5)
    assert got.size > 0 and same_bounds(gb, rb)
    differing = assert_cells_match(got, ref, shape)
    print(f"{shape} res={res} layers={layers}: {got.size} cells, {differing} normal components differ in the last bit")
    assert differing <= max(3, got.size // 1000)
    # no normals given: every cell keeps the GridCell default and the records are identical byte for byte
    got0, _ = ctx.generate_cells(area, None, res, layers, 1.5)
    ref0, _ = O.generate_cells(area, None, res, layers, 1.5)
    assert np.array_equal(got0.view(np.uint8), ref0.view(np.uint8))
    # the device and host implementations of the product agree on the kept set too
    host, hb = ctxmod.host_generate_cells(area, normals, res, layers, 1.5)
    assert same_bounds(gb, hb) and np.array_equal(host["x"], got["x"]) and np.array_equal(host["z"], got["z"])


def test_cell_lattice_awkward_inputs(ctx):
    rng = np.random.default_rng(5)
    area, normals = scenes.excavation_area("rect", 0.05)
    area = area.copy()
    normals = normals.astype(np.float32).copy()
    # non-finite points (dropped by the index, NaN skipped by the bounds), NaN normals (skipped by the sum)
    bad = rng.choice(area.shape[0], 40, replace=False)
    area[bad[:20], rng.integers(0, 3, 20)] = np.nan
    normals[bad[20:]] = np.nan
    ref, rb = O.generate_cells(area, normals, 0.1, 7, 1.5)
    got, gb = ctx.generate_cells(area, normals, 0.1, 7, 1.5)
    assert same_bounds(gb, rb)
    assert_cells_match(got, ref, "nan")
    # PointXYZRGB-like records with unaligned field offsets: point_step 22, x/y/z at 1/7/13
    n = area.shape[0]
    rec = np.zeros((n, 22), np.uint8)
    raw = area.view(np.uint8).reshape(n, 12)
    rec[:, 1:5], rec[:, 7:11], rec[:, 13:17] = raw[:, 0:4], raw[:, 4:8], raw[:, 8:12]
    got2, gb2 = ctx.generate_cells(rec.reshape(-1), normals, 0.1, 7, 1.5, point_step=22, off=(1, 7, 13))
    assert same_bounds(gb2, rb) and np.array_equal(got2.view(np.uint8), got.view(np.uint8))
    # far from the origin (float32 spacing 2^-11 m at 8 km) and a coarse lattice
    far = (area + np.float32([8000.0, -4000.0, 120.0])).astype(np.float32)
    ok = np.isfinite(far).all(1)
    ref3, rb3 = O.generate_cells(far[ok], None, 0.25, 4, 1.5)
    got3, gb3 = ctx.generate_cells(far[ok], None, 0.25, 4, 1.5)
    assert same_bounds(gb3, rb3) and np.array_equal(got3.view(np.uint8), ref3.view(np.uint8)) and got3.size > 0
    # empty area, no layers, capacity too small
    e, _ = ctx.generate_cells(np.zeros((0, 3), np.float32), None, 0.1, 10, 1.5)
    assert e.size == 0
    z, zb = ctx.generate_cells(area, None, 0.1, 0, 1.5)
    assert z.size == 0 and same_bounds(zb, O.generate_cells(area, None, 0.1, 0, 1.5)[1])
    v = A.cloud_view(area)
    import ctypes as C
    n_out = C.c_uint64(0)
    one = np.zeros(1, A.CELL_DTYPE)
    st = A.lib().plc_vl_generate_cells(ctx._h, C.byref(v), None, 0.1, 7, 1.5, one.ctypes.data, 1, C.byref(n_out), None)
    assert st == A.PLC_E_CAPACITY and n_out.value == ref.size and one["x"][0] == 0.0


def test_candidates_match_oracle(ctx, scene_cache):
    s = scene_cache("small")
    ix = O.Index(s.terrain, O.GRID)
    ctx.set_terrain(s.terrain)
    total = 0
    for num, height, radius in ((36, 1.1, 3.0), (100, 1.5, 3.0), (7, 1.1, 3.0), (1024, 1.1, 5.0), (4096, 2.3, 8.0)):
        ref = O.generate_candidates(ix, s.bounds, s.fixed, num, radius, height)
        for terrain in (s.terrain, None):        # the caller's cloud, or the one plc_vl_set_terrain left on the device
            got = ctx.generate_candidates(terrain, s.bounds, s.fixed, num, radius, height)
            assert got.size == ref.size > 0
            for k in ("x", "y", "z"):
                assert np.array_equal(got[k].view(np.uint64), ref[k].view(np.uint64)), (num, k)
            for k in ("pitch", "yaw"):
                assert np.allclose(got[k], ref[k], rtol=0, atol=4e-15), (num, k)
        total += ref.size
        # the ground was found under most poses (z is not just the sensor height)
        assert np.count_nonzero(ref["z"] != height) > 0
    assert total > 3000
    with pytest.raises(A.PlcError):              # a lattice of side 1 divides by zero upstream (:563)
        ctx.generate_candidates(s.terrain, s.bounds, s.fixed, 1, 3.0, 1.1)


def test_candidates_without_terrain_and_with_awkward_clouds(ctx, scene_cache):
    s = scene_cache("small")
    # empty terrain: getGroundHeight answers 0 everywhere (:601)
    empty = np.zeros((0, 3), np.float32)
    got = ctx.generate_candidates(empty, s.bounds, s.fixed, 64, 3.0, 1.1)
    ref = O.generate_candidates(O.Index(empty, O.GRID), s.bounds, s.fixed, 64, 3.0, 1.1)
    assert got.size == ref.size > 0 and np.array_equal(got["z"], ref["z"]) and np.all(got["z"] == 1.1)
    # NaN points, points far above the 2 m ball, a point exactly under a lattice node
    rng = np.random.default_rng(9)
    t = s.terrain[:: 7].copy()
    t[rng.choice(t.shape[0], 50, replace=False), 2] = np.nan
    t[rng.choice(t.shape[0], 500, replace=False), 2] += np.float32(3.0)
    ref = O.generate_candidates(O.Index(t, O.GRID), s.bounds, s.fixed, 400, 3.0, 1.1)
    got = ctx.generate_candidates(t, s.bounds, s.fixed, 400, 3.0, 1.1)
    assert got.size == ref.size
    for k in ("x", "y", "z"):
        assert np.array_equal(got[k].view(np.uint64), ref[k].view(np.uint64)), k


def test_pca_normals_close_to_oracle(ctx):
    for shape, radius in (("rect", 0.3), ("L", 1.5)):
        area, analytic = scenes.excavation_area(shape, 0.05)
        area = area.copy()
        area[11] = np.nan                          # a non-finite point gets a NaN normal and is nobody's neighbour
        got = ctx.compute_normals(area, radius)
        ref = O.compute_normals(area, radius)
        assert got.shape == ref.shape and np.isnan(got[11]).all()
        assert np.array_equal(np.isfinite(got).all(1), np.isfinite(ref).all(1))
        ok = np.isfinite(ref).all(1)
        assert ok.mean() > 0.99
        dots = np.abs(np.sum(got[ok].astype(np.float64) * ref[ok].astype(np.float64), 1))
        # same plane fit up to eigen-solver round-off (degenerate neighbourhoods aside)
        assert np.percentile(dots, 1) > 0.999999, np.percentile(dots, 1)
        assert np.all(got[ok][:, 2] >= 0)
        host = ctxmod.host_compute_normals(area, radius)
        hd = np.abs(np.sum(got[ok].astype(np.float64) * host[ok].astype(np.float64), 1))
        assert np.percentile(hd, 1) > 0.999999
    # an isolated pair has fewer than three neighbours: NaN (PCL)
    lonely = np.array([[0, 0, 0], [0.1, 0, 0], [50, 50, 50]], np.float32)
    assert np.isnan(ctx.compute_normals(lonely, 0.5)).all()


def test_full_size_lattices_have_the_reference_properties(ctx, scene_cache):
    """BASELINE config sizes, where the oracle is slow: properties instead of a twin run."""
    s = scene_cache("cfg2")
    area, normals = scenes.excavation_area("rect", 0.05)
    cells, b = ctx.generate_cells(area, normals, 0.01, 40, 1.5)       # 1.5e6 lattice points
    assert cells.size > 10_000            # the ball of 1.5 res is smaller than the cloud's 5 cm pitch: a sparse shell
    # lattice order: y index, x index, layer (lexicographic in (y, x, z))
    key = np.stack([cells["y"], cells["x"], cells["z"]], 1)
    assert np.all((np.diff(key[:, 0]) > 0) | ((np.diff(key[:, 0]) == 0) & ((np.diff(key[:, 1]) > 0) | ((np.diff(key[:, 1]) == 0) & (np.diff(key[:, 2]) > 0)))))
    n = normals_of(cells).astype(np.float64)
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-6) and np.all(cells["nz"] >= 0)
    # every kept cell is within 1.5 res of an area point and a sample of dropped lattice points is not
    ix = O.Index(area, O.GRID)
    rng = np.random.default_rng(3)
    for j in rng.choice(cells.size, 300, replace=False):
        assert ix.count((np.float32(cells["x"][j]), np.float32(cells["y"][j]), np.float32(cells["z"][j])), 0.01 * 1.5) > 0
    # candidates at config 3's lattice: 64 x 64, device-resident terrain, and the host version as the twin
    ctx.set_terrain(s.terrain)
    got = ctx.generate_candidates(None, s.bounds, s.fixed, 64 * 64, 3.0, 1.1)
    host = ctxmod.host_generate_candidates(s.terrain, s.bounds, s.fixed, 64 * 64, 3.0, 1.1)
    assert got.size == host.size > 3000
    for k in ("x", "y", "z"):
        assert np.array_equal(got[k], host[k]), k
    assert np.allclose(got["pitch"], host["pitch"], rtol=0, atol=4e-15)


def test_random_stage_scenes_against_oracle(ctx):
    """tools/fuzz_stages.py, a few iterations: random clouds, parameters, transforms and wire layouts through the
    generator, the setup steps and the drivable grid."""
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))
    import fuzz_stages
    tot = fuzz_stages.run(ctx, 10, 20260105)
    print(tot)
    assert tot["clouds"] == 10 and tot["cells"] > 0 and tot["poses"] > 0
