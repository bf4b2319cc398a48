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
