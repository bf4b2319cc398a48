import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_processor_b200 import Context, scenes
name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
sc = scenes.build(name)
with Context(0) as ctx:
    def t(f, n=10):
        f(); ctx.sync()
        t0 = time.perf_counter()
        for _ in range(n): f()
        ctx.sync()
        return (time.perf_counter() - t0) / n * 1e3
    out = dict(scene=name)
    out["set_terrain_ms"] = t(lambda: ctx.set_terrain(sc.terrain))
    out["set_cells_ms"] = t(lambda: ctx.set_cells(sc.cells))
    out["sweep_host_ms"] = t(lambda: ctx.sweep(sc.params, sc.fixed, sc.candidates, scores=True))
    out["kernel_ms"] = ctx.last_sweep_kernel_ms()
    ctx.upload_candidates(sc.params, sc.fixed, sc.candidates)
    out["run_async_sync_ms"] = t(lambda: (ctx.run_async(False, False), ctx.sync()))
    print(json.dumps(out))
