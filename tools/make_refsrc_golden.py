"""Golden vectors from the reference's OWN source: runs oracle/_ref/virtual_lidar_ref — /root/reference/src/virtual_lidar.cpp,
unmodified, compiled against stand-in PCL / ROS2 headers (`make -C oracle ref`) — on seeded scenes and stores what the
node publishes and logs in tests/golden/refsrc_<name>.npz, next to the inputs.  Needs /root/reference (this container);
the fixtures travel, the binary's sources do not.

    python tools/make_refsrc_golden.py            # rewrites every fixture
"""
import json
import os
import re
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, os.fspath(ROOT))
from pointcloud_processor_b200 import scenes  # noqa: E402

EXE = ROOT / "oracle" / "_ref" / "virtual_lidar_ref"
# name -> (pit shape, terrain points, terrain seed, num_candidates, vertical_layers, base transform xyz, observed cloud stride or 0)
CASES = {
    "L": ("L", 20_000, 21, 49, 10, (0.25, -0.5, 0.0), 0),
    "rect": ("rect", 30_000, 7, 100, 6, (0.0, 0.0, 0.1), 0),
    "L_observed": ("L", 20_000, 21, 36, 10, (0.25, -0.5, 0.0), 9),
}


def records(xyz):
    """PointXYZRGB wire records (point_step 32, x/y/z at 0/4/8, rgb at 16), as the upstream publishers send."""
    r = np.zeros((xyz.shape[0], 8), np.float32)
    r[:, :3] = xyz
    return r


def run_reference(area, terrain, observed, num_candidates, layers, base):
    subprocess.check_call(["make", "-C", os.fspath(ROOT / "oracle"), "ref"], stdout=subprocess.DEVNULL)
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        lines = [f"param num_candidates int {num_candidates}", f"param vertical_layers int {layers}"]
        for topic, cloud in (("/excavation_area", area), ("/excavated_terrain", terrain), ("/zx120/filtered_points", observed)):
            if cloud is None:
                continue
            f = tmp / (topic.strip("/").replace("/", "_") + ".bin")
            cloud.tofile(f)
            lines.append(f"cloud {topic} {f} {cloud.shape[0]} 32 map 10 20 x:0 y:4 z:8 rgb:16")
        lines += ["tf map zx120/base_link " + " ".join(repr(float(v)) for v in base) + " 0 0 0 1", "timers"]
        (tmp / "script.txt").write_text("\n".join(lines) + "\n")
        env = dict(os.environ, MOCK_ROS_SCRIPT=os.fspath(tmp / "script.txt"), MOCK_ROS_OUT=os.fspath(tmp))
        subprocess.run([os.fspath(EXE)], env=env, check=True, capture_output=True)
        msgs = []
        for j in sorted(tmp.glob("[0-9][0-9][0-9].json")):
            b = j.with_suffix(".bin")
            msgs.append((json.loads(j.read_text()), b.read_bytes() if b.exists() else b""))
        log = (tmp / "log.txt").read_text()
    return msgs, log


def parse(msgs, log):
    rows = lambda b: np.frombuffer(b, np.float64).reshape(-1, 14)
    by_topic = {}
    for m, b in msgs:
        by_topic.setdefault(m["topic"], []).append((m, b))
    grid = rows(by_topic["/excavation_grid_visualization"][-1][1])[1:]          # after the sweep; row 0 is DELETEALL
    cand = rows(by_topic["/mobile_lidar_candidate_positions"][-1][1])
    best = by_topic["/optimal_mobile_lidar_position"][-1][0]["point"]
    num = lambda pat: [float(x) for x in re.findall(pat, log)]
    counts = {k: [int(v) for v in re.findall(rf"{k} \([^)]*\): (\d+) cells", log)] for k in ("Green", "Red", "Blue", "Yellow")}
    return dict(cells_xyz=grid[:, 3:6].copy(), cell_rgb=grid[:, 9:12].copy(), cands_xyz=cand[2:-1, 3:6].copy(),
                fixed_xyz=cand[1, 3:6].copy(), best_xyz=np.array(best), fixed_total_2dp=num(r"Total Score \(ZX120 only\): ([-\d.]+)")[-1],
                best_score_2dp=num(r"\nINFO Total Score: ([-\d.]+)")[-1],
                colour_counts=np.array([[counts[k][i] for k in ("Green", "Red", "Blue", "Yellow")] for i in (0, 1)]),   # [zx120 only, dual]
                log=log)


def main():
    if not Path("/root/reference/src/virtual_lidar.cpp").exists():
        raise SystemExit("needs /root/reference")
    for name, (shape, n_pts, seed, num, layers, base, obs_stride) in CASES.items():
        terrain = records(scenes.terrain(shape, n_pts, seed=seed, side=24.0))
        area = records(scenes.excavation_area(shape, 0.05)[0])
        observed = records(area[::obs_stride, :3].copy()) if obs_stride else None
        out = parse(*run_reference(area, terrain, observed, num, layers, base))
        path = ROOT / "tests" / "golden" / f"refsrc_{name}.npz"
        np.savez_compressed(path, area=area[:, :3].copy(), terrain=terrain[:, :3].copy(),
                            observed=observed[:, :3].copy() if observed is not None else np.zeros((0, 3), np.float32),
                            num_candidates=num, vertical_layers=layers, base=np.array(base), **out)
        print(name, "cells", out["cells_xyz"].shape[0], "candidates", out["cands_xyz"].shape[0], "best", out["best_xyz"],
              "score", out["best_score_2dp"], "fixed", out["fixed_total_2dp"], "colours", out["colour_counts"].tolist(),
              "->", path.relative_to(ROOT), f"{path.stat().st_size / 1e3:.0f} kB")


if __name__ == "__main__":
    main()
