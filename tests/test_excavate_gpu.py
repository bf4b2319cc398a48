"""The excavated_surface_generator node's data path on the device (plc_excavate) against the oracle's
restatement of excavated_surface_generator.cpp:183-584, through the C ABI.

Bars: the kept input points (set, order, bytes incl. colour), every generated x / y, the colours and the output
order are bit-identical.  A generated z is `mean height - constant`; the mean is a double sum of hundreds of
float z values that upstream runs in KD-tree result order and the kernel in lane/tree order, so z may differ in
the last float32 bit (tolerance 1 ulp, count printed)."""
import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import _abi as A

pytestmark = pytest.mark.gpu


def ground(n, seed, lo=(-4.0, -7.0), hi=(12.0, 9.0), z0=0.0, rgb=True):
    """An uncarved /matched_point_cloud: PointXYZRGB records (n, 8) float32 with packed colours at byte 16."""
    rng = np.random.default_rng(seed)
    xy = rng.uniform(lo, hi, (n, 2))
    z = z0 + 0.15 * np.sin(0.4 * xy[:, 0]) * np.cos(0.3 * xy[:, 1]) + rng.uniform(-0.01, 0.01, n)
    rec = np.zeros((n, 8), np.float32)
    rec[:, :2], rec[:, 2], rec[:, 3] = xy, z, 1.0
    colours = (rng.integers(0, 1 << 24, n).astype(np.uint32) | np.uint32(0xFF000000)) if rgb else np.full(n, 0xFF000000, np.uint32)
    rec.view(np.uint32)[:, 4] = colours
    return rec, colours


def quat_z(yaw):
    return np.array([0.0, 0.0, np.sin(yaw / 2), np.cos(yaw / 2)])


def fields(records):
    f = records.view(np.float32).reshape(-1, 8)
    u = records.view(np.uint32).reshape(-1, 8)
    return f[:, 0], f[:, 1], f[:, 2], u[:, 3], u[:, 4], u[:, 5:]


def ulp_diff(a, b):
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    return np.abs(ia - ib)


def assert_clouds_match(got, ref, n_exact, what):
    """The first n_exact records byte-identical (kept input points); the rest identical except z within 1 ulp."""
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    assert np.array_equal(got[:n_exact], ref[:n_exact]), what
    gx, gy, gz, gw, gc, gp = fields(got[n_exact:])
    rx, ry, rz, rw, rc, rp = fields(ref[n_exact:])
    assert np.array_equal(gx.view(np.uint32), rx.view(np.uint32)) and np.array_equal(gy.view(np.uint32), ry.view(np.uint32)), what
    assert np.array_equal(gw, rw) and np.array_equal(gc, rc) and np.array_equal(gp, rp), what
    d = ulp_diff(gz, rz)
    assert d.max(initial=0) <= 1, (what, int(d.max()))
    return int(np.count_nonzero(d))


def run_both(ctx, rec, colours, t, q, prm):
    ns, na = A.excavation_counts(prm)
    ref_t, ref_a, ri = O.excavate(rec[:, :3], colours, t, q, prm, ns, na)
    got_t, got_a, gi = ctx.excavate(rec, t, q, prm)
    assert (gi.n_kept, gi.n_surface, gi.n_area) == (ri.n_kept, ri.n_surface, ri.n_area) == (ri.n_kept, ns, na)
    assert list(gi.center)[:2] == list(ri.center)[:2] and gi.yaw == ri.yaw
    assert abs(gi.center[2] - ri.center[2]) <= 1e-12 * max(1.0, abs(ri.center[2]))
    dz_t = assert_clouds_match(got_t, ref_t, int(ri.n_kept), "terrain")
    dz_a = assert_clouds_match(got_a, ref_a, 0, "area")
    return got_t, got_a, gi, dz_t + dz_a


@pytest.mark.parametrize("name,kw,n,t,yaw", [
    ("L default", {}, 60_000, (0.0, 0.0, 0.0), 0.0),
    ("rectangle", dict(l_shape_enabled=0), 60_000, (0.0, 0.0, 0.0), 0.0),
    ("L turned", dict(point_density=0.08, terrain_search_radius=0.3, depth=0.7, slope_angle_deg=60.0), 40_000, (1.5, -0.5, 0.2), 0.7),
    ("rectangle turned", dict(l_shape_enabled=0, length=2.4, width=0.9, offset_x=3.0, offset_y=-1.0), 40_000, (-0.5, 1.0, 0.0), -2.2),
    # excavated_surface_generator_L.cpp: wall points of the area cloud straight above the edge, offsets (5, 0)
    ("L file variant", dict(area_walls_vertical=1, offset_x=5.0, offset_y=0.0), 40_000, (0.0, 0.0, 0.0), 0.4),
])
def test_generator_matches_oracle(ctx, name, kw, n, t, yaw):
    prm = A.excavation_params(**kw)
    rec, colours = ground(n, 31)
    got_t, got_a, info, dz = run_both(ctx, rec, colours, np.array(t), quat_z(yaw), prm)
    print(f"{name}: {n} in, {info.n_kept} kept + {info.n_surface} surface, area {info.n_area}; {dz} z values differ in the last bit")
    assert 0 < n - info.n_kept < n // 4               # the pit was carved out of the cloud
    assert dz <= max(5, (info.n_surface + info.n_area) // 200)
    # colours: bottoms yellow / slopes darker yellow in the area cloud; dark green / light green in the terrain
    ac = fields(got_a)[4]
    assert set(np.unique(ac)) == {0xFFFFFF00, 0xFFC8C800}
    tc = fields(got_t[info.n_kept:])[4]
    assert set(np.unique(tc)) == {0xFF008B00, 0xFF90EE90}


def test_reference_default_sizes(ctx):
    # SURVEY §A.6 (measured on the reference's own rule): L area cloud 5 575 points of which 1 680 bottom; rectangle 3 130 / 888
    for kw, n_area, n_bottom in ((dict(), 5575, 1680), (dict(l_shape_enabled=0), 3130, 888)):
        prm = A.excavation_params(**kw)
        rec, _ = ground(30_000, 5)
        rec[:, 2] = 0.0
        _, area, info = ctx.excavate(rec, np.zeros(3), quat_z(0.0), prm)
        assert area.shape[0] == info.n_area == n_area
        assert int(np.count_nonzero(fields(area)[4] == 0xFFFFFF00)) == n_bottom
        assert np.all(fields(area)[2][fields(area)[4] == 0xFFFFFF00] == -1.0)      # flat ground at 0: bottoms at -depth


def test_heights_through_the_nearest_neighbour_fallback(ctx):
    # terrain 3 m above z = 0: no point lies in the 0.5 m ball around (x, y, 0), so every height is the z of the
    # nearest point in 3-D (:219-223)
    prm = A.excavation_params(point_density=0.1)
    rec, colours = ground(6_000, 7, lo=(1.0, -3.0), hi=(8.0, 4.0), z0=3.0)
    got_t, got_a, info, dz = run_both(ctx, rec, colours, np.zeros(3), quat_z(0.0), prm)
    assert dz == 0                                    # no sums involved: bit-identical
    # a sparse cloud far from the pit: queries walk many rings
    far = rec.copy()
    far[:, 0] += 40.0
    run_both(ctx, far, colours, np.zeros(3), quat_z(0.3), prm)


def test_awkward_inputs(ctx):
    prm = A.excavation_params()
    ns, na = A.excavation_counts(prm)
    # empty cloud: getTerrainHeight answers 0 and the synthetic surfaces are still produced (:185)
    empty = np.zeros((0, 8), np.float32)
    t, a, info = ctx.excavate(empty, np.zeros(3), quat_z(0.0), prm)
    rt, ra, _ = O.excavate(np.zeros((0, 3), np.float32), None, np.zeros(3), quat_z(0.0), prm, ns, na)
    assert np.array_equal(t, rt) and np.array_equal(a, ra) and t.shape[0] == ns and a.shape[0] == na
    # non-finite points fail every comparison of the carve test and are kept; no rgb field: a = 255, r = g = b = 0
    rec, _ = ground(20_000, 9)
    rng = np.random.default_rng(2)
    rec[rng.choice(rec.shape[0], 30, replace=False), rng.integers(0, 3, 30)] = np.nan
    xyz = np.ascontiguousarray(rec[:, :3])
    got_t, got_a, gi = ctx.excavate(xyz, np.zeros(3), quat_z(0.0), prm)
    ref_t, ref_a, ri = O.excavate(xyz, None, np.zeros(3), quat_z(0.0), prm, ns, na)
    assert gi.n_kept == ri.n_kept
    assert_clouds_match(got_t, ref_t, int(ri.n_kept), "nan terrain")
    assert_clouds_match(got_a, ref_a, 0, "nan area")
    assert np.all(fields(got_t[: gi.n_kept])[4] == 0xFF000000)
    # capacity too small: counts are reported, nothing is written
    import ctypes as C
    v = A.cloud_view(xyz)
    nt, nar = C.c_uint64(0), C.c_uint64(0)
    small = np.zeros((4, 32), np.uint8)
    tq = np.zeros(3), quat_z(0.0)
    st = A.lib().plc_excavate(ctx._h, C.byref(v), 0xFFFFFFFF, tq[0].ctypes.data, tq[1].ctypes.data, C.byref(prm),
                              small.ctypes.data, 4, C.byref(nt), small.ctypes.data, 4, C.byref(nar), None)
    assert st == A.PLC_E_CAPACITY and nt.value == ref_t.shape[0] and nar.value == na and not small.any()


def test_full_size_cloud_properties(ctx):
    """Config-4 scale (2 M merged points), where the oracle's per-point KD queries take minutes: properties."""
    prm = A.excavation_params()
    rec, colours = ground(2_000_000, 11, lo=(-11.0, -14.0), hi=(19.0, 16.0))
    terrain, area, info = ctx.excavate(rec, np.zeros(3), quat_z(0.0), prm)
    ns, na = A.excavation_counts(prm)
    assert info.n_surface == ns and info.n_area == na and terrain.shape[0] == info.n_kept + ns
    # kept points are a subsequence of the input, bytes untouched
    x, y, z, w, c, _ = fields(terrain[: info.n_kept])
    # carved points lie inside the widened outline and within [-depth, 0] of the local height
    inx = np.isin(rec[:, 0].view(np.uint32).astype(np.uint64) << 32 | rec[:, 1].view(np.uint32), x.view(np.uint32).astype(np.uint64) << 32 | y.view(np.uint32))
    carved = rec[~inx]
    assert carved.shape[0] == rec.shape[0] - info.n_kept > 1000
    grow = 1.0 / np.tan(np.deg2rad(75.0)) + 1e-6
    xl, yl = carved[:, 0] - 4.0, carved[:, 1] - 1.0
    in1 = (np.abs(xl) <= 0.6 + grow) & (np.abs(yl + 1.0) <= 1.0 + grow)
    in2 = (np.abs(xl - 1.0) <= 1.0 + grow) & (np.abs(yl + 1.4) <= 0.6 + grow)
    assert np.all(in1 | in2)
    # order preserved: the kept x sequence equals the input's with the carved rows dropped
    assert np.array_equal(x.view(np.uint32), rec[inx, 0].view(np.uint32)) and np.array_equal(c, colours[inx])
    # a sample of kept points inside the outline is checked one by one against the oracle's height rule
    inside = np.flatnonzero(inx & (np.abs(rec[:, 0] - 4.0) < 0.5) & (np.abs(rec[:, 1]) < 0.5))
    assert inside.size > 0       # points above the local mean height (z_relative > 0) survive inside the pit
