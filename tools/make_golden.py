"""Writes the golden fixtures under tests/golden/ (run here, on CPU; commit the outputs).

The reference ships no golden vectors and cannot be executed ("parity unpinned"),
so the fixtures are of two kinds:

  hand_cases.json   scenes small enough to derive the answer by hand from the reference's source
                    (virtual_lidar.cpp:754-800 and the PCL radiusSearch contract): the expected
                    values in the file were written down from that reasoning, NOT produced by
                    running any code, and both the oracle and the CUDA path are tested against them.
  *.npz             seeded inputs together with the outputs of the oracle's BRUTE-FORCE index
                    (O(N) scan per radius search, no acceleration structure, one thread).  They pin
                    the oracle's accelerated indices, the OpenMP path and the CUDA path to one
                    simple, auditable evaluation of the restated algorithm, and catch regressions.

    python tools/make_golden.py
"""
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, os.fspath(ROOT))
from oracle import binding as O  # noqa: E402
from pointcloud_processor_b200 import _abi as A, scenes  # noqa: E402

OUT = ROOT / "tests" / "golden"


def params_dict(p):
    return {k: getattr(p, k) for k, _ in A.Params._fields_}


def sweep_fixture(name, scene, observed=None, params=None, n_cands=None):
    prm = params or scene.params
    cands = scene.candidates if n_cands is None else scene.candidates[:n_cands]
    ref = O.sweep(O.Index(scene.terrain, O.BRUTE), None if observed is None else O.Index(observed, O.BRUTE),
                  scene.cells, prm, scene.fixed, cands, threads=1)
    np.savez_compressed(
        OUT / f"{name}.npz", terrain=scene.terrain, observed=np.zeros((0, 3), np.float32) if observed is None else observed,
        cells=scene.cells, fixed=scene.fixed, candidates=cands,
        params=np.array(json.dumps(params_dict(prm))),
        mask=ref["mask"], scores=ref["scores"], covered=ref["covered"], flags=ref["flags"],
        fixed_scores=ref["fixed_scores"], best_index=ref["best_index"], best_score=ref["best_score"],
        fixed_total=ref["fixed_total"], samples=ref["counters"]["samples"],
        fixed_samples=ref["counters"]["fixed_samples"], rays_visible=ref["counters"]["rays_visible"])
    print(name, "C", cands.size, "M", scene.cells.size, "N", scene.terrain.shape[0], "best", ref["best_index"],
          "visible", ref["counters"]["rays_visible"], "samples", ref["counters"]["samples"])


def filter_merge_fixture():
    rng = np.random.default_rng(77)
    n = 6000
    scan = rng.uniform([-3, -12, -2.5], [18, 12, 11], (n, 3)).astype(np.float32)
    scan[::97, 0] = np.nan
    scan[5::211, 2] = np.inf
    scan[3] = (0.0, 0.0, 0.0)          # x == 0: strict bound, dropped
    scan[4] = (15.0, 1.0, 1.0)         # x == front: dropped
    scan[6] = (1.0, -10.0, 1.0)        # y == -side: dropped
    scan[7] = (1.0, 1.0, -1.5)         # z == -1.5: dropped
    box = A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0)
    cropped, idx = O.crop(scan, box)
    cent, keys, counts, pk = O.voxel_grid(cropped, 0.2)
    # launch/tf_four_wheel_robot.launch.xml:4 (0.25 0 2.0, pitch 0.1245) composed with a map pose
    t = np.array([12.5, -3.25, 0.75])
    half = 0.1245 / 2
    q = np.array([0.0, np.sin(half), 0.0, np.cos(half)])
    moved = O.transform(cent, t, q)
    np.savez_compressed(OUT / "filter_merge.npz", scan=scan, box=np.array([0.0, 15.0, -10.0, 10.0, -1.5, 10.0]),
                        leaf=0.2, crop_index=idx, cropped=cropped, centroids=cent, voxel_keys=keys,
                        voxel_counts=counts, point_keys=pk, t=t, q=q, transformed=moved)
    print("filter_merge", "n", n, "cropped", cropped.shape[0], "voxels", cent.shape[0])


def hand_cases():
    """Known answers derived by hand from virtual_lidar.cpp:754-800 (see each `why`)."""
    r = 0.08 * 0.7   # :782
    cases = [
        dict(name="point_on_second_sample",
             sensor=[0, 0, 0], pitch=0.0, cell=[3, 0, 0], normal=[0, 0, 1], terrain=[[0.8, 0, 0]], visible=False, samples=2,
             why="samples at s=0.5, 0.8, 1.1, ...; the terrain point sits exactly on the sample at 0.8 => d2 = 0 < r^2, "
                 "blocked at the 2nd sample"),
        dict(name="point_between_samples_tunnels",
             sensor=[0, 0, 0], pitch=0.0, cell=[3, 0, 0], normal=[0, 0, 1], terrain=[[0.65, 0, 0]], visible=True, samples=9,
             why="0.65 is 0.15 m from the samples at 0.5 and 0.8, r = 0.056 => no sample sees it (SURVEY F3); "
                 "end = 3 - 0.08 = 2.92, samples 0.5..2.9 => 9 samples"),
        dict(name="point_just_inside_radius",
             sensor=[0, 0, 0], pitch=0.0, cell=[3, 0, 0], normal=[0, 0, 1], terrain=[[0.8, r - 1e-4, 0]], visible=False, samples=2,
             why="lateral offset r - 1e-4 < r => d2 < r^2 at the sample at 0.8"),
        dict(name="point_just_outside_radius",
             sensor=[0, 0, 0], pitch=0.0, cell=[3, 0, 0], normal=[0, 0, 1], terrain=[[0.8, r + 1e-4, 0]], visible=True, samples=9,
             why="lateral offset r + 1e-4 > r => the strict d2 < r^2 test fails at every sample"),
        dict(name="point_near_target_is_ignored",
             sensor=[0, 0, 0], pitch=0.0, cell=[3, 0, 0], normal=[0, 0, 1], terrain=[[2.99, 0, 0]], visible=True, samples=9,
             why="the loop stops at s < L - 0.08 = 2.92; the last sample is at 2.9, 0.09 m from the point > r"),
        dict(name="short_ray_has_no_samples",
             sensor=[0, 0, 0], pitch=0.0, cell=[0.55, 0, 0], normal=[0, 0, 1], terrain=[[0.3, 0, 0]], visible=True, samples=0,
             why="end = 0.55 - 0.08 = 0.47 < start 0.5 => the while loop never runs (:769), so the ray is visible"),
        dict(name="out_of_range_scores_zero",
             sensor=[0, 0, 0], pitch=0.0, cell=[0.4, 0, 0], normal=[0, 0, 1], terrain=[[5, 5, 5]], visible=False, samples=0,
             why="L = 0.4 < MIN_DISTANCE 0.5 => evaluateCellScore returns 0 before any visibility test (:662-669)"),
        dict(name="outside_vertical_fov",
             sensor=[0, 0, 0], pitch=-2.0, cell=[3, 0, 0], normal=[0, 0, 1], terrain=[[5, 5, 5]], visible=False, samples=0,
             why="elevation 0, pitch -2.0 rad: |0 - (-2.0)| = 2.0 > pi/2 => out of FOV (:703-714)"),
    ]
    # score of the visible horizontal ray onto a cell with normal +z: incidence term |b.n| = 0 => score = 1/L
    for c in cases:
        L = float(np.linalg.norm(np.array(c["cell"], float) - np.array(c["sensor"], float)))
        c["score"] = (1.0 / L) if c["visible"] else 0.0
    (OUT / "hand_cases.json").write_text(json.dumps(dict(
        about="hand-derived known answers for the sampled-sphere ray test; see tools/make_golden.py",
        cases=cases), indent=1))
    print("hand_cases", len(cases))


def node_stage_fixtures():
    """The nodes either side of the sweep: seeded inputs and the oracle's outputs.
    Radius searches go through the BRUTE index where the oracle lets the caller choose one."""
    rng = np.random.default_rng(2024)
    # N2: a ground cloud around the pit -> /excavated_terrain + /excavation_area (coarser lattice to keep the file small)
    n = 9000
    xy = rng.uniform([0.5, -3.5], [8.5, 3.5], (n, 2))
    z = 0.15 * np.sin(0.4 * xy[:, 0]) * np.cos(0.3 * xy[:, 1]) + rng.uniform(-0.01, 0.01, n)
    ground = np.column_stack([xy, z]).astype(np.float32)
    colours = (rng.integers(0, 1 << 24, n).astype(np.uint32) | np.uint32(0xFF000000))
    prm = A.excavation_params(point_density=0.1)
    ns, na = A.excavation_counts(prm)
    base_t, base_q = np.array([0.25, -0.5, 0.0]), np.array([0.0, 0.0, np.sin(0.15), np.cos(0.15)])
    terrain, area, info = O.excavate(ground, colours, base_t, base_q, prm, ns, na)
    exc = dict(ground=ground, colours=colours, base_t=base_t, base_q=base_q, point_density=0.1,
               terrain=terrain, area=area, n_kept=info.n_kept, center=np.array(list(info.center)), yaw=info.yaw)
    # N1 on the generator's outputs: normals, cells, candidates
    area_xyz = np.ascontiguousarray(area[:, :12]).view(np.float32).reshape(-1, 3)
    terrain_xyz = np.ascontiguousarray(terrain[:, :12]).view(np.float32).reshape(-1, 3)
    normals = O.compute_normals(area_xyz, 1.5)
    cells, b = O.generate_cells(area_xyz, normals, 0.2, 4, 1.5)
    fixed = scenes.fixed_pose()
    cands = O.generate_candidates(O.Index(terrain_xyz, O.BRUTE), b, fixed, 49, 3.0, 1.1)
    bounds = np.array([b.min_x, b.max_x, b.min_y, b.max_y, b.min_z, b.max_z, b.width, b.height], np.float64)
    # N4: a scan in the robot's sensor frame -> occupancy grid
    m = 8000
    sxy = rng.uniform(-12, 12, (m, 2))
    sz = -2.0 + 0.2 * np.sin(0.5 * sxy[:, 0]) + rng.normal(0, 0.02, m) + (rng.random(m) < 0.1) * rng.uniform(0, 2, m)
    scan = np.column_stack([sxy, sz]).astype(np.float32)
    scan[::131, 1] = np.nan
    dt, dq = np.array([9.25, -3.0, 2.0]), np.array([-0.06, 0.0166, 0.963, 0.2675])
    dq = dq / np.linalg.norm(dq)
    grid = O.drivable_area(scan, dt, dq, (9.0, -3.0), (7.0, -2.0), grid_resolution=0.5, map_width=30.0, map_height=24.0,
                           min_points_per_cell=3)
    np.savez_compressed(OUT / "node_stages.npz", normals=normals, cells=cells, bounds=bounds, fixed=fixed, candidates=cands,
                        scan=scan, drivable_t=dt, drivable_q=dq, grid=grid, **exc)
    print("node_stages", "ground", n, "kept", info.n_kept, "surface", ns, "area", na, "cells", cells.size, "cands", cands.size,
          "grid", grid.shape, {int(v): int((grid == v).sum()) for v in np.unique(grid)})


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    hand_cases()
    node_stage_fixtures()
    sweep_fixture("sweep_tiny", scenes.build("tiny"))
    small = scenes.build("small")
    sweep_fixture("sweep_small_first8", small, n_cands=8)
    # memo parameterisation + an observed cloud for the fixed sensor (virtual_lidar_memo.cpp)
    tiny = scenes.build("tiny")
    rng = np.random.default_rng(5)
    pick = rng.permutation(tiny.area.shape[0])[:400]
    observed = (tiny.area[pick] + rng.normal(0, 0.02, (pick.size, 3))).astype(np.float32)
    sweep_fixture("sweep_tiny_memo_observed", tiny, observed=observed, params=A.memo_params())
    sweep_fixture("sweep_tiny_default_observed", tiny, observed=observed)
    filter_merge_fixture()


if __name__ == "__main__":
    main()
