"""Wall time of each C-ABI call of one end-to-end step (pinned host buffers), next to the library's kernel time.

    python tools/dev/e2e_phases.py [cfg2]
"""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from pointcloud_processor_b200 import scenes  # noqa: E402
from pointcloud_processor_b200.context import Context  # noqa: E402

sc = scenes.build(sys.argv[1] if len(sys.argv) > 1 else "cfg2")
keep = []


def pinned(a):
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).pin_memory()
    keep.append(t)
    return t.numpy().view(a.dtype).reshape(a.shape)


terrain, cells, cands = pinned(sc.terrain), pinned(sc.cells), pinned(sc.candidates)
ctx = Context(0)
walls = {"set_terrain": [], "set_cells": [], "sweep": []}
kern = {"terrain": [], "sweep": []}
for rep in range(12):
    t0 = time.perf_counter()
    ctx.set_terrain(terrain)
    t1 = time.perf_counter()
    ctx.set_cells(cells)
    t2 = time.perf_counter()
    ctx.sweep(sc.params, sc.fixed, cands, scores=True)
    t3 = time.perf_counter()
    walls["set_terrain"].append((t1 - t0) * 1e3)
    walls["set_cells"].append((t2 - t1) * 1e3)
    walls["sweep"].append((t3 - t2) * 1e3)
    kern["sweep"].append(ctx.last_sweep_kernel_ms())
    kern["terrain"].append(ctx.device_times())
out = {k: float(np.median(v[2:])) for k, v in walls.items()}
out["k_sweep_ms"] = float(np.median(kern["sweep"][2:]))
out["device_times_last"] = kern["terrain"][-1]
print(json.dumps(out))
