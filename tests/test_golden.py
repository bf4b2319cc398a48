"""The oracle against the committed golden fixtures (CPU).  tests/golden/*.npz hold the outputs of
the oracle's brute-force index (O(N) scan, one thread) on seeded inputs; hand_cases.json holds
answers derived by hand from the reference source.  The reference itself ships no vectors
(parity unpinned), so these are what pins the oracle."""
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
