"""The CUDA path (through the C ABI) against the committed golden fixtures in tests/golden/."""
import numpy as np
import pytest

from golden_util import SWEEP_FIXTURES, hand_cases, hand_scene, load_sweep, GOLDEN
from pointcloud_processor_b200 import _abi as A

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", hand_cases(), ids=lambda c: c["name"])
def test_hand_derived_answers(ctx, case):
    terrain, cells, pose, fixed = hand_scene(case)
    ctx.set_terrain(terrain)
    ctx.set_observed(np.zeros((0, 3), np.float32))
    ctx.set_cells(cells)
    got = ctx.sweep(A.default_params(), fixed, pose, mask=True)
    assert bool(got.mask[0, 0] & 1) == case["visible"], case["why"]
    assert got.counters["samples"] == case["samples"], case["why"]
    assert got.scores[0] == pytest.approx(case["score"], rel=1e-12, abs=1e-15)


@pytest.mark.parametrize("name", SWEEP_FIXTURES)
def test_sweep_fixture(ctx, name):
    z, prm = load_sweep(name)
    ctx.set_terrain(z["terrain"])
    ctx.set_observed(z["observed"])
    ctx.set_cells(z["cells"])
    got = ctx.sweep(prm, z["fixed"], z["candidates"], mask=True, flags=True, fixed_scores=True)
    assert np.array_equal(got.mask, z["mask"])                       # bit-exact visibility
    np.testing.assert_allclose(got.scores, z["scores"], rtol=1e-12, atol=1e-12)
    assert np.array_equal(got.covered, z["covered"])
    assert got.best_index == int(z["best_index"])
    np.testing.assert_allclose(got.fixed_scores, z["fixed_scores"], rtol=1e-12, atol=1e-15)
    assert np.array_equal(got.cell_flags, z["flags"])
    assert got.counters["samples"] == int(z["samples"])
    assert got.counters["fixed_samples"] == int(z["fixed_samples"])
    assert got.counters["rays_visible"] == int(z["rays_visible"])


def test_filter_merge_fixture(ctx):
    z = np.load(GOLDEN / "filter_merge.npz")
    box = A.CropBox(*map(float, z["box"]))
    out, n_crop = ctx.filter(z["scan"], box, float(z["leaf"]))
    assert n_crop == z["cropped"].shape[0]
    assert np.array_equal(out[:, :3].view(np.uint32), z["centroids"].view(np.uint32))
    merged, each = ctx.merge([dict(cloud=np.ascontiguousarray(out[:, :3]), t=z["t"], q=z["q"], rgb=(255, 0, 0))])
    xyz = merged[:, :12].copy().view(np.float32).reshape(-1, 3)
    np.testing.assert_allclose(xyz, z["transformed"], rtol=1e-6, atol=1e-6)
    assert np.array_equal(xyz.view(np.uint32), z["transformed"].view(np.uint32))


def test_node_stage_fixture(ctx):
    """The device versions of the generator, the setup steps and the drivable grid against tests/golden/node_stages.npz."""
    from test_excavate_gpu import assert_clouds_match
    from test_golden import bounds_of
    z = np.load(GOLDEN / "node_stages.npz")
    prm = A.excavation_params(point_density=float(z["point_density"]))
    n = z["ground"].shape[0]
    rec = np.zeros((n, 8), np.float32)
    rec[:, :3], rec[:, 3] = z["ground"], 1.0
    rec.view(np.uint32)[:, 4] = z["colours"]
    terrain, area, info = ctx.excavate(rec, z["base_t"], z["base_q"], prm)
    assert info.n_kept == int(z["n_kept"]) and info.yaw == float(z["yaw"])
    assert_clouds_match(terrain, z["terrain"], int(z["n_kept"]), "terrain")
    assert_clouds_match(area, z["area"], 0, "area")
    area_xyz = np.ascontiguousarray(z["area"][:, :12]).view(np.float32).reshape(-1, 3)
    terrain_xyz = np.ascontiguousarray(z["terrain"][:, :12]).view(np.float32).reshape(-1, 3)
    cells, b = ctx.generate_cells(area_xyz, z["normals"], 0.2, 4, 1.5)
    assert cells.size == z["cells"].size
    for k in ("x", "y", "z"):
        assert np.array_equal(cells[k], z["cells"][k])
    for k in ("nx", "ny", "nz"):
        assert np.max(np.abs(cells[k] - z["cells"][k])) <= 1.2e-7
    assert [b.min_x, b.max_x, b.min_y, b.max_y, b.min_z, b.max_z, b.width, b.height] == list(z["bounds"])
    cands = ctx.generate_candidates(terrain_xyz, bounds_of(z), z["fixed"], 49, 3.0, 1.1)
    assert cands.size == z["candidates"].size
    for k in ("x", "y", "z"):
        assert np.array_equal(cands[k], z["candidates"][k])
    for k in ("pitch", "yaw"):
        assert np.allclose(cands[k], z["candidates"][k], rtol=0, atol=4e-15)
    normals = ctx.compute_normals(area_xyz, 1.5)
    ok = np.isfinite(z["normals"]).all(1)
    assert np.array_equal(ok, np.isfinite(normals).all(1))
    assert np.percentile(np.abs(np.sum(normals[ok].astype(np.float64) * z["normals"][ok], 1)), 1) > 0.99999
    prm_d = A.DrivableParams()
    A.lib().plc_drivable_params_default(prm_d)
    prm_d.grid_resolution, prm_d.map_width, prm_d.map_height, prm_d.min_points_per_cell = 0.5, 30.0, 24.0, 3
    grid, _ = ctx.drivable_area(z["scan"], z["drivable_t"], z["drivable_q"], (9.0, -3.0), (7.0, -2.0), prm_d)
    assert np.array_equal(grid, z["grid"])
