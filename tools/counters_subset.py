"""k_sweep on a subset of a scene's poses (for profiling the large configs): python tools/counters_subset.py cfg5 4000"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_processor_b200 import Context, scenes
name, n = sys.argv[1], int(sys.argv[2])
sc = scenes.build(name)
sel = np.sort(np.random.default_rng(3).choice(sc.candidates.size, min(n, sc.candidates.size), replace=False))
cands = sc.candidates[sel]
with Context(0) as ctx:
    ctx.set_terrain(sc.terrain); ctx.set_cells(sc.cells)
    if os.environ.get("PLC_COUNT"): ctx.count_samples(True)     # counting mode instead of the default any-hit
    for _ in range(2):
        r = ctx.sweep(sc.params, sc.fixed, cands, scores=False)
    g = ctx.grid_info()[0]
    print(json.dumps(dict(scene=name, poses=int(cands.size), kernel_ms=ctx.last_sweep_kernel_ms(), **r.counters, n_vox=g["n_voxels"], dim=list(g["dim"]))))
