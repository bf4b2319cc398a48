import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_processor_b200 import Context, scenes
for name in sys.argv[1:]:
    sc = scenes.build(name)
    with Context(0) as ctx:
        ctx.set_terrain(sc.terrain); ctx.set_cells(sc.cells)
        if os.environ.get("PLC_COUNT"): ctx.count_samples(True)     # counting mode instead of the default any-hit
        r = ctx.sweep(sc.params, sc.fixed, sc.candidates[:8], scores=False)      # builds the index the sweep wants
        build_ms = ctx.device_times()
        r = ctx.sweep(sc.params, sc.fixed, sc.candidates, scores=False)
        t0=time.time(); r = ctx.sweep(sc.params, sc.fixed, sc.candidates, scores=False); dt=time.time()-t0
        g = ctx.grid_info()[0]
        print(json.dumps(dict(scene=name, ms=dt*1e3, kernel_ms=ctx.last_sweep_kernel_ms(), **r.counters, grid_bytes=g["grid_bytes"], n_vox=g["n_voxels"], set_terrain_ms=build_ms.get("set_terrain_ms"))))
