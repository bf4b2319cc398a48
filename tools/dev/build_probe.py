"""Warm device time of plc_vl_set_terrain (K1) per scene: python tools/dev/build_probe.py cfg2 cfg3"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from pointcloud_processor_b200 import Context, scenes
for name in sys.argv[1:]:
    sc = scenes.build(name) if name != "cfg5" else None
    terr = sc.terrain if sc else scenes.terrain("L", 4_000_000, 6, 40.0)
    with Context(0) as ctx:
        ks, ws = [], []
        for _ in range(6):
            t0 = time.perf_counter(); ctx.set_terrain(terr); ws.append((time.perf_counter() - t0) * 1e3)
            ks.append(ctx.device_times()["set_terrain_ms"])
        g = ctx.grid_info()[0]
        print(json.dumps(dict(scene=name, refine=os.environ.get("PLC_DEV_REFINE", "1"), kernel_ms=float(np.median(ks[1:])),
                              wall_ms=float(np.median(ws[1:])), n_vox=g["n_voxels"])))
