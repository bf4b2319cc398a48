"""Randomised differential test of plc_filter / plc_merge against the oracle (run on the GPU box):
random scans (sizes across the block and tile boundaries, NaN/inf coordinates, dense clumps that fill single
voxels), record layouts (point_step 12..48, unaligned offsets), crop boxes and leaf sizes, random rigid transforms.

    python tools/fuzz_filter.py [iterations] [seed]
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import binding as O  # noqa: E402
from pointcloud_processor_b200 import Context, _abi as A  # noqa: E402


def scene(rng):
    n = int(rng.choice([0, 1, 2, 255, 256, 257, 1000, 4097, 30000, 120000]))
    span = float(rng.choice([1.0, 20.0, 80.0]))
    xyz = rng.uniform(-span, span, (n, 3))
    if n and rng.random() < 0.5:                      # dense clumps: many points per voxel, long segments
        k = int(rng.integers(1, 6))
        centres = rng.uniform(0, 10, (k, 3))
        idx = rng.integers(0, k, n)
        xyz = centres[idx] + rng.normal(0, float(rng.choice([0.01, 0.1, 1.0])), (n, 3))
    xyz = xyz.astype(np.float32)
    if n > 10 and rng.random() < 0.4:
        xyz[rng.integers(0, n, 5), rng.integers(0, 3, 5)] = np.nan
        xyz[rng.integers(0, n, 3), rng.integers(0, 3, 3)] = np.inf
    step = int(rng.choice([12, 16, 20, 32, 48, 15]))
    offs = (0, 4, 8) if rng.random() < 0.7 else tuple(int(o) for o in sorted(rng.choice(np.arange(0, step - 3), 3, replace=False)))
    if offs != (0, 4, 8) and (offs[1] - offs[0] < 4 or offs[2] - offs[1] < 4):
        offs = (0, 4, 8)
    if step < 12 or offs[2] + 4 > step:
        step, offs = 16, (0, 4, 8)
    buf = np.zeros(n * step, np.uint8)
    rec = buf.reshape(n, step) if n else buf.reshape(0, step)
    for a in range(3):
        rec[:, offs[a]:offs[a] + 4] = xyz[:, a:a + 1].copy().view(np.uint8)
    lo = rng.uniform(-span, 0, 3)
    hi = lo + rng.uniform(0.5, 2 * span, 3)
    box = A.CropBox(lo[0], hi[0], lo[1], hi[1], lo[2], hi[2]) if rng.random() < 0.7 else A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0)
    leaf = float(rng.choice([0.2, 0.05, 1.0, 0.0, 3.7, 0.001]))
    return xyz, buf, step, offs, box, leaf


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    bad = 0
    pts_in = vox = 0
    t0 = time.time()
    with Context(0) as ctx:
        for it in range(iters):
            rng = np.random.default_rng(seed * 7919 + it)
            xyz, buf, step, offs, box, leaf = scene(rng)
            out, n_crop = ctx.filter(buf, box, leaf, point_step=step, off=offs)
            cxyz, _ = O.crop(xyz, box)
            want = cxyz
            if leaf > 0 and cxyz.shape[0]:
                vg = O.voxel_grid(cxyz, leaf)
                if vg is not None:
                    want = vg[0]
            ok = n_crop == cxyz.shape[0] and out.shape[0] == want.shape[0] and \
                np.array_equal(out[:, :3].view(np.uint32), want.view(np.uint32))
            pts_in += xyz.shape[0]
            vox += want.shape[0]
            if ok and out.shape[0]:
                q = rng.normal(size=4)
                q /= np.linalg.norm(q)
                t = rng.uniform(-50, 50, 3)
                merged, each = ctx.merge([dict(cloud=out, t=t, q=q, rgb=(255, 0, 0), present=1),
                                          dict(cloud=out[: out.shape[0] // 2].copy(), t=-t, q=q[::-1].copy(), rgb=(0, 0, 255), present=1)])
                m = merged.view(np.float32).reshape(-1, 8)[:, :3]
                e0 = O.transform(out[:, :3], t, q)
                e1 = O.transform(out[: out.shape[0] // 2, :3], -t, q[::-1].copy())
                ok = np.array_equal(m[: e0.shape[0]].view(np.uint32), e0.view(np.uint32)) and \
                    np.array_equal(m[e0.shape[0]:].view(np.uint32), e1.view(np.uint32))
            if not ok:
                bad += 1
                print("MISMATCH it", it, "seed", seed, "n", xyz.shape[0], "step", step, offs, "leaf", leaf, "crop", n_crop,
                      cxyz.shape[0], "out", out.shape[0], want.shape[0], flush=True)
    print(f"fuzz_filter: {iters} scans, {bad} mismatching, {pts_in} points in, {vox} voxels out, {time.time() - t0:.1f} s")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
