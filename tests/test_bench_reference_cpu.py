"""bench.py's reference arm on the host alone (it never touches the GPU or libplc.so): the informational legs that time
the reference's own sources (oracle/_ref) and SURVEY D5's feeder baseline run, agree with each other, and degrade to an
`unavailable` note where oracle/_ref was not built."""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from pointcloud_processor_b200 import scenes  # noqa: E402

HAVE_REF = (ROOT / "oracle" / "_ref" / "virtual_lidar_ref").exists()


def test_feeders_cpu_baseline_port_and_reference_source_agree():
    r = bench.feeders_cpu_reference(n_scan=60_000)
    assert "unavailable" not in r, r
    assert r["filter_robot_port_ms"] > 0 and r["filter_zx120_port_ms"] > 0 and r["merge_port_ms"] > 0
    n_in, n_crop, n_vox = r["filter_robot_points"]
    assert n_in == 60_000 and 0 < n_vox < n_crop < n_in and r["merge_points"] == n_vox + r["filter_zx120_points"][2]
    if HAVE_REF:
        assert r["filter_robot_reference_source_voxels"] == n_vox          # the reference's own node keeps the same voxels
        assert r["filter_robot_reference_source_ms"] > 0


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref is built where the reference's sources are present")
def test_reference_source_sample_times_the_reference_node():
    sc = scenes.build("ref_default", setup=bench.oracle_setup())          # 50 k points, 3 814 cells: a second of sweep to difference
    r = bench.reference_source_sample(sc, sc.meta["res"], sc.meta["layers_found"], k_small=4, k_large=100)
    assert "unavailable" not in r, r
    assert r["kind"] == "reference" and r["cores"] == 1 and r["value"] > 1e4 and r["unit"] == bench.UNIT


def test_reference_source_sample_reports_a_missing_binary(monkeypatch, tmp_path):
    monkeypatch.setattr(bench, "ROOT", tmp_path)
    assert "unavailable" in bench.reference_source_sample(None, 0.1, 6)
