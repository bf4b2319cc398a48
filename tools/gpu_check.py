"""Developer check run on the GPU box: CUDA path vs CPU oracle on a few scenes, with timings.
Not part of the product; the pytest suite in tests/ is the judged parity gate."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import binding as O  # noqa: E402
from pointcloud_processor_b200 import Context, _abi as A, scenes  # noqa: E402


def check_scene(name, ctx, threads=8, full=True):
    t0 = time.time()
    sc = scenes.build(name)
    t_scene = time.time() - t0
    t0 = time.time()
    ctx.set_terrain(sc.terrain)
    ctx.set_cells(sc.cells)
    t_set = time.time() - t0
    t0 = time.time()
    got = ctx.sweep(sc.params, sc.fixed, sc.candidates, mask=full, flags=full, fixed_scores=True)
    t_gpu = time.time() - t0
    t0 = time.time()
    got2 = ctx.sweep(sc.params, sc.fixed, sc.candidates)
    t_gpu2 = time.time() - t0
    out = dict(scene=name, C=int(sc.candidates.size), M=int(sc.cells.size), N=int(sc.terrain.shape[0]),
               t_scene=round(t_scene, 3), t_set=round(t_set, 4), t_gpu_first=round(t_gpu, 4),
               t_gpu_second=round(t_gpu2, 4), counters=got.counters, best=got.best_index,
               rays_per_s=got.counters["rays"] / t_gpu2, grid=ctx.grid_info()[0])
    if full:
        ix = O.Index(sc.terrain, O.GRID)
        t0 = time.time()
        ref = O.sweep(ix, None, sc.cells, sc.params, sc.fixed, sc.candidates, threads=threads, want_mask=True)
        out["t_oracle_mt"] = round(time.time() - t0, 3)
        mism = int(np.count_nonzero(np.unpackbits((got.mask ^ ref["mask"]).view(np.uint8))))
        out["mask_mismatch"] = mism
        out["best_equal"] = got.best_index == ref["best_index"]
        out["score_max_rel"] = float(np.max(np.abs(got.scores - ref["scores"]) / np.maximum(1.0, np.abs(ref["scores"]))))
        out["covered_equal"] = bool(np.array_equal(got.covered, ref["covered"]))
        out["samples_equal"] = got.counters["samples"] == ref["counters"]["samples"]
        out["fixed_samples_equal"] = got.counters["fixed_samples"] == ref["counters"]["fixed_samples"]
        out["fixed_scores_max_abs"] = float(np.max(np.abs(got.fixed_scores - ref["fixed_scores"])))
        if sc.candidates.size <= 300:
            ref1 = O.sweep(ix, None, sc.cells, sc.params, sc.fixed, sc.candidates, threads=1, want_mask=False)
            out["flags_equal"] = bool(np.array_equal(got.cell_flags, ref1["flags"]))
    print(json.dumps(out), flush=True)
    return out


def check_filter_merge(ctx):
    rng = np.random.default_rng(5)
    n = 300_000
    pts = np.zeros((n, 8), np.float32)   # point_step 32, like a velodyne driver
    pts[:, :3] = rng.uniform([-5, -14, -3], [22, 14, 12], (n, 3)).astype(np.float32)
    pts[::1000, 1] = np.nan
    box = A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0)
    t0 = time.time()
    out, n_crop = ctx.filter(pts, box, 0.2)
    t1 = time.time() - t0
    cxyz, cidx = O.crop(pts, box)
    vg = O.voxel_grid(cxyz, 0.2)
    ok_crop = n_crop == cxyz.shape[0]
    ok_n = out.shape[0] == vg[0].shape[0]
    ok_bits = ok_n and np.array_equal(out[:, :3].view(np.uint32), vg[0].view(np.uint32))
    crop_only, _ = ctx.filter(pts, box, 0.0)
    ok_crop_bits = np.array_equal(crop_only[:, :3].view(np.uint32), cxyz.view(np.uint32))
    # merge
    q = np.array([0.0, 0.2164396, 0.0, 0.976296])   # pitch 0.4363
    t = np.array([0.55, 0.4, 3.5])
    srcs = [dict(cloud=out, t=[1.0, 2.0, 0.5], q=[0, 0, 0.38268343, 0.92387953], rgb=(255, 0, 0)),
            dict(cloud=crop_only[:5000].copy(), t=t, q=q, rgb=(0, 0, 255))]
    merged, each = ctx.merge(srcs)
    exp0 = O.transform(out, srcs[0]["t"], srcs[0]["q"])
    exp1 = O.transform(srcs[1]["cloud"], t, q)
    m = merged.view(np.float32).reshape(-1, 8)
    ok_merge = np.array_equal(m[: exp0.shape[0], :3].view(np.uint32), exp0.view(np.uint32)) and \
        np.array_equal(m[exp0.shape[0]:, :3].view(np.uint32), exp1.view(np.uint32))
    rgba = merged.view(np.uint32).reshape(-1, 8)[:, 4]
    ok_rgb = bool(np.all(rgba[: exp0.shape[0]] == 0xFFFF0000) and np.all(rgba[exp0.shape[0]:] == 0xFF0000FF))
    print(json.dumps(dict(filter_ms=round(t1 * 1e3, 2), n_crop=n_crop, n_vox=int(out.shape[0]), ok_crop=bool(ok_crop),
                          ok_n=bool(ok_n), ok_centroid_bits=bool(ok_bits), ok_crop_bits=bool(ok_crop_bits),
                          ok_merge_bits=bool(ok_merge), ok_rgb=ok_rgb, each=[int(e.shape[0]) for e in each])), flush=True)


if __name__ == "__main__":
    names = sys.argv[1:] or ["tiny", "small", "ref_default", "cfg1"]
    with Context(0) as ctx:
        ctx.count_samples(True)          # the evaluated-sample counter is compared below
        check_filter_merge(ctx)
        for nm in names:
            check_scene(nm, ctx, full=nm not in ("cfg3", "cfg5"))
