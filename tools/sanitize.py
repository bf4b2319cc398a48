"""Small end-to-end run of every kernel of the library on awkward inputs (partial tiles, out-of-range poses,
empty clouds, scenes far from the origin), for a build with bounds checks on the occupancy and list accesses:

    make -C pointcloud_processor_b200/csrc DEV_FLAGS=-DPLC_BOUNDS_CHECK ..., then
    PLC_DEV_LIB=.../dev_chk.so python tools/sanitize.py      -> prints the out-of-range access count (must be 0)

(compute-sanitizer is closed on the GPU pool: its wrapper there refuses to run and says so.)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_processor_b200 import Context, _abi as A, scenes  # noqa: E402
from pointcloud_processor_b200.nodes import run_pipeline  # noqa: E402

with Context(0) as ctx:
  for counting in (True, False):       # both traversal modes of k_sweep
    ctx.count_samples(counting)
    for name in ("tiny", "small", "cfg1"):
        sc = scenes.build(name)
        ctx.set_terrain(sc.terrain)
        ctx.set_cells(sc.cells)
        ctx.set_observed(sc.area[::7].copy())
        r = ctx.sweep(sc.params, sc.fixed, sc.candidates, mask=True, flags=True, fixed_scores=True)
        print(name, r.best_index, r.counters)
        prm = A.default_params()
        prm.fov_vertical, prm.max_distance, prm.min_distance = 0.3, 6.0, 2.0
        far = sc.candidates.copy()
        far["x"] += 1000.0
        for c in (sc.candidates, far, sc.candidates[:1], sc.candidates[:0]):
            r = ctx.sweep(prm, sc.fixed, c, mask=True)
        ctx.set_terrain(np.zeros((0, 3), np.float32))
        try:
            ctx.sweep(sc.params, sc.fixed, sc.candidates)
            raise SystemExit("a sweep without terrain must be refused")
        except A.PlcError:
            pass
        ctx.set_terrain(sc.terrain + np.float32(5000.0))
        r = ctx.sweep(sc.params, sc.fixed, sc.candidates)
    inp = scenes.pipeline_inputs(n_scan=20_000, sweep="tiny")
    got = run_pipeline(ctx, inp, mask=True)
    print("pipeline", got["merged"].shape, got["sweep"].best_index)
    L = A.lib()
    if hasattr(L, "plc_dev_oob_count"):
        import ctypes
        L.plc_dev_oob_count.restype = ctypes.c_ulonglong
        n = L.plc_dev_oob_count()
        print("out-of-range accesses:", n)
        assert n == 0
print("done")
