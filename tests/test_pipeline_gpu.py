"""BASELINE config 4: two 32-ring scans -> crop -> voxel downsample -> merge -> sweep.
Reduced sizes are compared stage by stage with the oracle; the full-size run is checked through
size-independent properties."""
import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import Context, _abi as A, scenes
from pointcloud_processor_b200.nodes import run_pipeline

pytestmark = pytest.mark.gpu

BOX = A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0)      # pointcloud_filter.cpp:30-36, :106-116


def xyz(rec):
    return rec[:, :12].copy().view(np.float32).reshape(-1, 3)


def rgb(rec):
    return rec[:, 16:20].copy().view(np.uint32).ravel() & 0xFFFFFF


def test_pipeline_stages_match_oracle(ctx):
    inp = scenes.pipeline_inputs(n_scan=60_000, sweep="small")
    got = run_pipeline(ctx, inp, mask=True, flags=True, fixed_scores=True)
    sc = inp.scene
    # filter: identical voxel membership and order, centroids bit for bit
    exp = {}
    for name, scan in (("robot", inp.robot_scan), ("zx120", inp.zx120_scan)):
        cropped, _ = O.crop(scan, BOX)
        cent = O.voxel_grid(cropped, 0.2)[0]
        exp[name] = cent
        out = got[f"{name}_filtered"]
        assert out.shape[0] == cent.shape[0] > 500
        assert np.array_equal(out[:, :3].view(np.uint32), cent.view(np.uint32))
    # merge: robot (red) first, then backhoe (blue), coordinates bit for bit
    nr, nb = exp["robot"].shape[0], exp["zx120"].shape[0]
    m = got["merged"]
    assert m.shape[0] == nr + nb
    assert np.array_equal(xyz(m)[:nr].view(np.uint32), O.transform(exp["robot"], *inp.robot_tf).view(np.uint32))
    assert np.array_equal(xyz(m)[nr:].view(np.uint32), O.transform(exp["zx120"], *inp.zx120_tf).view(np.uint32))
    assert np.all(rgb(m)[:nr] == 0xFF0000) and np.all(rgb(m)[nr:] == 0x0000FF)
    assert np.array_equal(got["robot_coloured"], m[:nr]) and np.array_equal(got["backhoe_coloured"], m[nr:])
    # sweep with the zx120 filter output as the fixed sensor's observed cloud
    ref = O.sweep(O.Index(sc.terrain, O.GRID), O.Index(exp["zx120"], O.GRID), sc.cells, sc.params, sc.fixed,
                  sc.candidates, threads=4, want_mask=True)
    r = got["sweep"]
    assert np.array_equal(r.mask, ref["mask"])
    assert r.best_index == ref["best_index"]
    assert np.allclose(r.scores, ref["scores"], rtol=1e-12, atol=0)
    assert np.array_equal(r.covered, ref["covered"])
    assert np.max(np.abs(r.fixed_scores - ref["fixed_scores"])) <= 1e-12
    assert r.counters["samples"] == ref["counters"]["samples"]
    assert all(v > 0 for v in got["device_ms"].values())


def test_pipeline_full_size_properties(ctx):
    inp = scenes.pipeline_inputs(n_scan=1_000_000, sweep="cfg1")
    got = run_pipeline(ctx, inp)
    for name, scan in (("robot", inp.robot_scan), ("zx120", inp.zx120_scan)):
        out = got[f"{name}_filtered"]
        p = scan[:, :3].astype(np.float64)
        keep = (p[:, 0] > 0) & (p[:, 0] < 15) & (np.abs(p[:, 1]) < 10) & (p[:, 2] > -1.5) & (p[:, 2] < 10)
        kept = scan[keep, :3]
        # one centroid per occupied voxel, ascending PCL index, each inside its own voxel
        inv = np.float32(1.0) / np.float32(0.2)
        ijk = np.floor(kept * inv).astype(np.int64)
        mn = ijk.min(0)
        div = ijk.max(0) - mn + 1
        key = (ijk[:, 0] - mn[0]) + (ijk[:, 1] - mn[1]) * div[0] + (ijk[:, 2] - mn[2]) * div[0] * div[1]
        uniq = np.unique(key)
        assert out.shape[0] == uniq.size
        oijk = np.floor(out[:, :3] * inv).astype(np.int64)
        okey = (oijk[:, 0] - mn[0]) + (oijk[:, 1] - mn[1]) * div[0] + (oijk[:, 2] - mn[2]) * div[0] * div[1]
        # a centroid on a voxel face can round across it; membership is by construction, so allow a handful
        assert np.count_nonzero(okey != uniq) <= max(3, uniq.size // 2000)
        # filtering is a projection on the crop: cropping the output again changes nothing
        again = (out[:, 0] > 0) & (out[:, 0] < 15) & (np.abs(out[:, 1]) < 10) & (out[:, 2] > -1.5) & (out[:, 2] < 10)
        assert again.all()
        # the centroid of centroids weighted by counts is the centroid of the cropped cloud (linearity)
        counts = np.bincount(np.searchsorted(uniq, key), minlength=uniq.size)
        lhs = (out[:, :3].astype(np.float64) * counts[:, None]).sum(0) / counts.sum()
        assert np.allclose(lhs, kept.astype(np.float64).mean(0), atol=1e-4)
    nr, nb = got["robot_filtered"].shape[0], got["zx120_filtered"].shape[0]
    m = got["merged"]
    assert m.shape[0] == nr + nb
    # rigid transform: pairwise distances survive (checked on a sample)
    a = got["zx120_filtered"][:2000, :3].astype(np.float64)
    b = xyz(m)[nr:nr + 2000].astype(np.float64)
    assert np.allclose(np.linalg.norm(a[1:] - a[:-1], axis=1), np.linalg.norm(b[1:] - b[:-1], axis=1), atol=1e-4)
    assert got["sweep"].best_index >= 0
