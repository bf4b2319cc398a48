"""Randomised differential test of the CUDA sweep against the oracle (run on the GPU box; not part of the suite):
random clouds, cells (with non-finite and duplicate entries, partial tiles), poses (inside the cloud, far away,
on top of a cell) and parameters (radii, steps, ranges, FOVs, both fixed-sensor modes, observed clouds).

    python tools/fuzz_sweep.py [iterations] [seed]
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import binding as O  # noqa: E402
from pointcloud_processor_b200 import Context, _abi as A  # noqa: E402


def scene(rng):
    scale = float(rng.choice([1.0, 1.0, 3.0, 0.4]))
    off = rng.choice([0.0, 0.0, 0.0, 250.0, -4000.0]) * np.array([1.0, -0.7, 0.05])
    n_floor, n_wall, n_dust = int(rng.integers(0, 6000)), int(rng.integers(0, 3000)), int(rng.integers(0, 500))
    floor = np.stack([rng.uniform(-4, 4, n_floor), rng.uniform(-4, 4, n_floor), rng.normal(0, 0.03, n_floor)], 1)
    wall = np.stack([rng.normal(1.5, 0.02, n_wall), rng.uniform(-2, 2, n_wall), rng.uniform(0, 1.5, n_wall)], 1)
    dust = rng.uniform([-4, -4, 0], [4, 4, 3], (n_dust, 3))
    terrain = (np.concatenate([floor, wall, dust]) * scale + off).astype(np.float32)
    if terrain.shape[0] == 0:
        terrain = (np.array([[0.3, 0.2, 0.1]]) * scale + off).astype(np.float32)
    if rng.random() < 0.2:
        terrain[rng.integers(0, terrain.shape[0], 5)] = np.nan
    m = int(rng.choice([1, 31, 33, 511, 512, 513, 700, 1500, 2600]))
    cells = np.zeros(m, A.CELL_DTYPE)
    pos = rng.uniform([-3, -3, 0.0], [3, 3, 1.4], (m, 3)) * scale + off
    if rng.random() < 0.5:                                    # a lattice, like the reference's cells
        g = np.stack(np.meshgrid(np.linspace(-3, 3, 30), np.linspace(-3, 3, 30), np.linspace(0.1, 1.3, 3), indexing="ij"), -1)
        pos = (g.reshape(-1, 3)[rng.permutation(2700)[:m]] if m <= 2700 else pos) * scale + off
        pos = pos[np.lexsort((pos[:, 2], pos[:, 1], pos[:, 0]))]
    cells["x"], cells["y"], cells["z"] = pos.T
    nrm = rng.normal(size=(m, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    cells["nx"], cells["ny"], cells["nz"] = nrm.astype(np.float32).T
    if rng.random() < 0.25 and m > 3:
        cells["x"][rng.integers(0, m)] = np.nan
        cells["z"][rng.integers(0, m)] = np.inf
        cells[rng.integers(0, m)] = cells[rng.integers(0, m)]
    c = int(rng.choice([2, 7, 20, 60]))
    poses = np.zeros(c, A.POSE_DTYPE)
    p = rng.uniform([-5, -5, 0.2], [5, 5, 5], (c, 3)) * scale + off
    if rng.random() < 0.3:
        p[0] = pos[rng.integers(0, m)]                       # a pose on top of a cell: L = 0
    if rng.random() < 0.3:
        p[-1] = p[-1] + 40.0 * scale                          # everything out of range
    poses["x"], poses["y"], poses["z"] = p.T
    # mostly looking at the cells' centre, with a spread that puts the FOV boundary across them
    centre = np.nanmean(np.where(np.isfinite(pos), pos, np.nan), axis=0)
    d = centre - p
    aim = np.arctan2(d[:, 2], np.hypot(d[:, 0], d[:, 1]))
    poses["pitch"] = np.where(rng.random(c) < 0.8, aim + rng.normal(0, 0.4, c), rng.uniform(-1.57, 0.6, c))
    poses["yaw"] = rng.uniform(-3.2, 3.2, c)
    fixed = np.zeros(1, A.POSE_DTYPE)
    fixed["x"], fixed["y"], fixed["z"] = off + np.array([0.4, 0.5, 3.5]) * scale
    fixed["pitch"] = -np.pi / 6
    prm = A.memo_params() if rng.random() < 0.25 else A.default_params()
    prm.ray_step = float(rng.choice([0.3, 0.2, 0.11, 0.5]))
    prm.vis_radius = float(rng.choice([0.08, 0.05, 0.15, 0.3]))
    prm.block_radius_scale = float(rng.choice([0.7, 1.0, 0.4]))
    prm.max_distance = float(rng.choice([15.0, 6.0, 3.0, 40.0])) * scale
    prm.min_distance = float(rng.choice([0.5, 0.5, 2.0, 0.0]))
    prm.fov_vertical = float(rng.choice([np.pi, 1.0, 0.3, 3.5]))
    observed = None
    if rng.random() < 0.3:
        observed = (pos[rng.integers(0, m, 300)] + rng.normal(0, 0.05, (300, 3))).astype(np.float32)
    return terrain, observed, cells, poses, fixed, prm


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    bad = 0
    tot = dict(rays=0, gated=0, visible=0, samples=0, exact=0)
    t0 = time.time()
    with Context(0) as ctx:
        ctx.count_samples(True)          # the evaluated-sample counter is compared below
        for it in range(iters):
            rng = np.random.default_rng(seed * 100003 + it)
            terrain, observed, cells, poses, fixed, prm = scene(rng)
            ctx.set_terrain(terrain)
            ctx.set_observed(observed if observed is not None else np.zeros((0, 3), np.float32))
            ctx.set_cells(cells)
            finite = terrain[np.isfinite(terrain).all(1)]
            if finite.shape[0] == 0:
                continue
            try:
                got = ctx.sweep(prm, fixed, poses, mask=True, flags=True, fixed_scores=True)
            except A.PlcError as e:
                print(it, "refused:", e)
                continue
            ref = O.sweep(O.Index(terrain, O.GRID), O.Index(observed, O.GRID) if observed is not None else None, cells, prm,
                          fixed, poses)
            mism = int(np.count_nonzero(np.unpackbits((got.mask ^ ref["mask"]).view(np.uint8))))
            with np.errstate(invalid="ignore"):
                rel = np.abs(got.scores - ref["scores"]) / np.maximum(1.0, np.abs(ref["scores"]))
            ok = (mism == 0 and np.nanmax(rel, initial=0.0) <= 1e-12 and np.array_equal(got.covered, ref["covered"])
                  and got.counters["samples"] == ref["counters"]["samples"]
                  and got.counters["rays_gated_in"] == ref["counters"]["rays_gated_in"]
                  and np.array_equal(got.cell_flags, ref["flags"])
                  and np.allclose(got.fixed_scores, ref["fixed_scores"], rtol=1e-12, atol=1e-15))
            # the any-hit traversal (the default of the library) on the same scene: every output but the sample counter
            ctx.count_samples(False)
            fast = ctx.sweep(prm, fixed, poses, mask=True, flags=True, fixed_scores=True)
            ctx.count_samples(True)
            ok = ok and (np.array_equal(fast.mask, got.mask) and np.array_equal(fast.covered, got.covered)
                         and np.array_equal(fast.scores, got.scores, equal_nan=True) and np.array_equal(fast.cell_flags, got.cell_flags)
                         and np.array_equal(fast.fixed_scores, got.fixed_scores, equal_nan=True) and fast.best_index == got.best_index
                         and fast.counters["rays_visible"] == got.counters["rays_visible"])
            tot["rays"] += got.counters["rays"]; tot["gated"] += got.counters["rays_gated_in"]
            tot["visible"] += got.counters["rays_visible"]; tot["samples"] += got.counters["samples"]
            tot["exact"] += got.counters["exact_tests"]
            if not ok:
                bad += 1
                print("MISMATCH it", it, "seed", seed, "mask", mism, "rel", float(np.nanmax(rel, initial=0.0)),
                      "samples", got.counters["samples"], ref["counters"]["samples"], "gated", got.counters["rays_gated_in"],
                      ref["counters"]["rays_gated_in"], "cells", cells.size, "poses", poses.size, flush=True)
    print(f"fuzz: {iters} scenes (seed {seed}) in both traversal modes, {bad} mismatching, {time.time() - t0:.1f} s, totals {tot}")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
