"""The oracle against the committed golden fixtures (CPU).  tests/golden/*.npz hold the outputs of
the oracle's brute-force index (O(N) scan, one thread) on seeded inputs; hand_cases.json holds
answers derived by hand from the reference source.  The reference itself ships no vectors
(SURVEY §8c), so these — beside the reference-source fixtures of test_reference_source.py — are what pins the oracle."""
import numpy as np
import pytest

from golden_util import SWEEP_FIXTURES, hand_cases, hand_scene, load_sweep, GOLDEN
from oracle import binding as O
from pointcloud_processor_b200 import _abi as A


@pytest.mark.parametrize("case", hand_cases(), ids=lambda c: c["name"])
@pytest.mark.parametrize("kind", [O.BRUTE, O.GRID, O.KDTREE])
def test_hand_derived_answers(case, kind):
    terrain, cells, pose, fixed = hand_scene(case)
    ref = O.sweep(O.Index(terrain, kind), None, cells, A.default_params(), fixed, pose)
    assert bool(ref["mask"][0, 0] & 1) == case["visible"], case["why"]
    assert ref["counters"]["samples"] == case["samples"], case["why"]
    assert ref["scores"][0] == pytest.approx(case["score"], rel=1e-12, abs=1e-15)


@pytest.mark.parametrize("name", SWEEP_FIXTURES)
@pytest.mark.parametrize("kind,threads", [(O.GRID, 1), (O.KDTREE, 1), (O.GRID, 4)])
def test_accelerated_oracle_reproduces_brute_force_fixture(name, kind, threads):
    z, prm = load_sweep(name)
    obs = O.Index(z["observed"], kind) if z["observed"].shape[0] else None
    ref = O.sweep(O.Index(z["terrain"], kind), obs, z["cells"], prm, z["fixed"], z["candidates"], threads=threads)
    assert np.array_equal(ref["mask"], z["mask"])
    assert np.array_equal(ref["scores"], z["scores"])          # same code path, same order: bit-identical
    assert np.array_equal(ref["covered"], z["covered"])
    assert ref["best_index"] == int(z["best_index"])
    assert np.array_equal(ref["fixed_scores"], z["fixed_scores"])
    assert ref["counters"]["samples"] == int(z["samples"])
    assert ref["counters"]["fixed_samples"] == int(z["fixed_samples"])
    if threads == 1:
        assert np.array_equal(ref["flags"], z["flags"])


def test_filter_merge_fixture():
    z = np.load(GOLDEN / "filter_merge.npz")
    box = A.CropBox(*map(float, z["box"]))
    cropped, idx = O.crop(z["scan"], box)
    assert np.array_equal(idx, z["crop_index"])
    assert np.array_equal(cropped.view(np.uint32), z["cropped"].view(np.uint32))
    # hand-checkable: the four points planted exactly on a bound are dropped (strict comparisons)
    assert not set([3, 4, 6, 7]) & set(idx.tolist())
    cent, keys, counts, pk = O.voxel_grid(cropped, float(z["leaf"]))
    assert np.array_equal(keys, z["voxel_keys"]) and np.array_equal(counts, z["voxel_counts"])
    assert np.array_equal(cent.view(np.uint32), z["centroids"].view(np.uint32))
    assert counts.sum() == cropped.shape[0] and np.all(np.diff(keys) > 0)
    moved = O.transform(cent, z["t"], z["q"])
    assert np.array_equal(moved.view(np.uint32), z["transformed"].view(np.uint32))


def bounds_of(z):
    b = A.GridBounds()
    (b.min_x, b.max_x, b.min_y, b.max_y, b.min_z, b.max_z), b.width, b.height = z["bounds"][:6], int(z["bounds"][6]), int(z["bounds"][7])
    return b


def test_node_stage_fixture():
    """tests/golden/node_stages.npz (tools/make_golden.py): generator -> setup -> drivable grid.  The oracle
    reproduces it through its other index kinds, and the product's host twins of the setup steps agree with it."""
    from pointcloud_processor_b200 import context as ctxmod
    z = np.load(GOLDEN / "node_stages.npz")
    prm = A.excavation_params(point_density=float(z["point_density"]))
    ns, na = A.excavation_counts(prm)
    terrain, area, info = O.excavate(z["ground"], z["colours"], z["base_t"], z["base_q"], prm, ns, na)
    assert np.array_equal(terrain, z["terrain"]) and np.array_equal(area, z["area"]) and info.n_kept == int(z["n_kept"])
    assert terrain.shape[0] == info.n_kept + ns and area.shape[0] == na
    area_xyz = np.ascontiguousarray(area[:, :12]).view(np.float32).reshape(-1, 3)
    terrain_xyz = np.ascontiguousarray(terrain[:, :12]).view(np.float32).reshape(-1, 3)
    cells, b = O.generate_cells(area_xyz, z["normals"], 0.2, 4, 1.5)
    assert np.array_equal(cells.view(np.uint8), z["cells"].view(np.uint8))
    hcells, hb = ctxmod.host_generate_cells(area_xyz, z["normals"], 0.2, 4, 1.5)
    assert np.array_equal(hcells.view(np.uint8), z["cells"].view(np.uint8))
    for kind in (O.GRID, O.KDTREE):
        cands = O.generate_candidates(O.Index(terrain_xyz, kind), bounds_of(z), z["fixed"], 49, 3.0, 1.1)
        assert np.array_equal(cands.view(np.uint8), z["candidates"].view(np.uint8))
    hc = ctxmod.host_generate_candidates(terrain_xyz, bounds_of(z), z["fixed"], 49, 3.0, 1.1)
    assert np.array_equal(hc.view(np.uint8), z["candidates"].view(np.uint8))
    grid = O.drivable_area(z["scan"], z["drivable_t"], z["drivable_q"], (9.0, -3.0), (7.0, -2.0), grid_resolution=0.5,
                           map_width=30.0, map_height=24.0, min_points_per_cell=3)
    assert np.array_equal(grid, z["grid"])
