"""Masks, scores and optimum of the library named by PLC_DEV_LIB against the oracle on cfg1 (full) — for k_sweep
variants whose sample counter is not the reference's (any-hit experiments).  python tools/dev/mask_check.py [scene]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from oracle import binding as O
from pointcloud_processor_b200 import Context, scenes
name = sys.argv[1] if len(sys.argv) > 1 else "cfg1"
s = scenes.build(name)
with Context(0) as ctx:
    ctx.set_terrain(s.terrain); ctx.set_observed(np.zeros((0, 3), np.float32)); ctx.set_cells(s.cells)
    got = ctx.sweep(s.params, s.fixed, s.candidates, mask=True, flags=True, fixed_scores=True)
ref = O.sweep(O.Index(s.terrain, O.GRID), None, s.cells, s.params, s.fixed, s.candidates)
mism = int(np.count_nonzero(np.unpackbits((got.mask ^ ref["mask"]).view(np.uint8))))
print("mask mismatches", mism, "max rel score err", float(np.max(np.abs(got.scores - ref["scores"]) / np.maximum(np.abs(ref["scores"]), 1e-300))),
      "argmax", got.best_index, ref["best_index"], "covered equal", bool(np.array_equal(got.covered, ref["covered"])),
      "samples", got.counters["samples"], ref["counters"]["samples"], "exact", got.counters.get("exact_tests"))
