"""Loaders for the committed fixtures in tests/golden/ (written by tools/make_golden.py)."""
import json
from pathlib import Path

import numpy as np

from pointcloud_processor_b200 import _abi as A

GOLDEN = Path(__file__).resolve().parent / "golden"
SWEEP_FIXTURES = ["sweep_tiny", "sweep_small_first8", "sweep_tiny_memo_observed", "sweep_tiny_default_observed"]


def load_sweep(name):
    z = np.load(GOLDEN / f"{name}.npz")
    prm = A.default_params()
    for k, v in json.loads(str(z["params"])).items():
        setattr(prm, k, v)
    return z, prm


def hand_cases():
    return json.loads((GOLDEN / "hand_cases.json").read_text())["cases"]


def hand_scene(case):
    """One mobile ray: terrain points, one cell, one pose; the fixed sensor sits out of range."""
    terrain = np.asarray(case["terrain"], np.float32).reshape(-1, 3)
    cells = np.zeros(1, A.CELL_DTYPE)
    cells["x"], cells["y"], cells["z"] = case["cell"]
    cells["nx"], cells["ny"], cells["nz"] = case["normal"]
    pose = np.zeros(1, A.POSE_DTYPE)
    pose["x"], pose["y"], pose["z"] = case["sensor"]
    pose["pitch"] = case["pitch"]
    fixed = np.zeros(1, A.POSE_DTYPE)
    fixed["x"] = 1000.0          # > max_distance from the cell: contributes score 0 and no samples
    return terrain, cells, pose, fixed
