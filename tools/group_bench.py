"""Single-process multi-GPU sweep through plc_group_* (what a ROS2 node would call): wall time per sweep.
    python tools/group_bench.py cfg3 8"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_processor_b200 import scenes
from pointcloud_processor_b200.context import Group

name, n_dev = sys.argv[1], int(sys.argv[2])
sc = scenes.build(name)
with Group(list(range(n_dev))) as g:
    t0 = time.perf_counter(); g.set_terrain(sc.terrain); t_terrain = time.perf_counter() - t0
    t0 = time.perf_counter(); g.set_cells(sc.cells); t_cells = time.perf_counter() - t0
    r = g.sweep(sc.params, sc.fixed, sc.candidates)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        r = g.sweep(sc.params, sc.fixed, sc.candidates)
        ts.append((time.perf_counter() - t0) * 1e3)
    print(json.dumps(dict(scene=name, devices=n_dev, poses=int(sc.candidates.size), rays_per_pose=int(sc.cells.size),
                          set_terrain_ms=t_terrain * 1e3, set_cells_ms=t_cells * 1e3, sweep_wall_ms=float(np.median(ts)),
                          rays_per_s=sc.candidates.size * sc.cells.size / (np.median(ts) * 1e-3),
                          best_index=r.best_index, best_score=r.best_score)))
