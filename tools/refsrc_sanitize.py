"""Plays the scripts behind tests/golden/refsrc_*.npz through the reference's own nodes built with AddressSanitizer and
UndefinedBehaviorSanitizer (`make -C oracle ref-asan`): a fixture must not be the product of undefined behaviour in the
reference's sources or in the stand-in headers.  Needs /root/reference (this container).  Exits non-zero on any report.

    python tools/refsrc_sanitize.py
"""
import os
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [os.fspath(ROOT), os.fspath(ROOT / "tools"), os.fspath(ROOT / "tests")]
import make_refsrc_golden as G  # noqa: E402
import refsrc_inputs as R  # noqa: E402
from pointcloud_processor_b200 import scenes  # noqa: E402


def run(node, make_lines, label):
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        (tmp / "script.txt").write_text("\n".join(make_lines(tmp)) + "\n")
        env = dict(os.environ, MOCK_ROS_SCRIPT=os.fspath(tmp / "script.txt"), MOCK_ROS_OUT=os.fspath(tmp), ASAN_OPTIONS="detect_leaks=0")
        r = subprocess.run([os.fspath(ROOT / "oracle" / "_build" / "ref_asan" / f"{node}_ref")], env=env, capture_output=True, text=True)
    print(f"{label:45s} rc {r.returncode}", r.stderr.strip()[-800:])
    return r.returncode != 0 or "runtime error" in r.stderr or "AddressSanitizer" in r.stderr


def main():
    subprocess.check_call(["make", "-C", os.fspath(ROOT / "oracle"), "ref-asan"], stdout=subprocess.DEVNULL)
    bad = 0
    C, V, E = R.cycles_inputs(scenes), R.vl_edges_inputs(scenes), R.edges_inputs(scenes)
    bad += run("virtual_lidar", lambda tmp: R.cycles_script(tmp, C, G.cloud_line, G.tf_line), "virtual_lidar: five cycles")
    bad += run("virtual_lidar", lambda tmp: R.vl_edges_script(tmp, V, G.cloud_line, G.tf_line), "virtual_lidar: guards")
    for name in ("pointcloud_filter", "calc_drivable_area", "excavated_surface_generator_disabled"):
        bad += run(name.replace("_disabled", ""), lambda tmp: R.edges_scripts(tmp, E, G.cloud_line, G.tf_line)[name], f"{name}: edge rules")
    f = np.load(ROOT / "tests" / "golden" / "refsrc_filter.npz")
    clouds = {w: f[f"{w}_records"] for w in ("robot", "backhoe")}
    bad += run("pointcloud_merger", lambda tmp: R.merger_script(tmp, clouds, G.cloud_line, G.tf_line), "pointcloud_merger")
    for name in ("pointcloud_filter", "calc_drivable_area", "excavated_surface_generator_L", "excavated_surface_generator_rect"):
        node = name if not name.startswith("excavated") else "excavated_surface_generator"
        bad += run(node, lambda tmp: R.variants_scripts(tmp, G.cloud_line, G.tf_line)[name], f"{name}: other parameters")
    print("reports:", bad)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
