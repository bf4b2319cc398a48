"""The whole dataflow of launch/start.launch.xml on the device (nodes.run_launch_graph): filter -> merger ->
terrain generator -> virtual_lidar setup -> sweep, with calc_drivable_area beside it.  Each stage is compared with
the oracle's restatement of the same node fed with the previous stage's output."""
import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import Context, _abi as A, scenes
from pointcloud_processor_b200.nodes import run_launch_graph, run_launch_graph_resident
from test_excavate_gpu import assert_clouds_match
from test_sweep_gpu import assert_same_optimum

pytestmark = pytest.mark.gpu

BOX = A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0)


def xyz(rec):
    return np.ascontiguousarray(rec[:, :12]).view(np.float32).reshape(-1, 3)


def test_launch_graph_stage_by_stage(ctx):
    inp = scenes.pipeline_inputs(n_scan=150_000, sweep="tiny")
    got = run_launch_graph(ctx, inp, num_candidates=64, mask=True, flags=True, fixed_scores=True)
    # filter + merger: bit for bit (covered in detail by test_pipeline_gpu)
    cent = {}
    for name, scan in (("robot", inp.robot_scan), ("zx120", inp.zx120_scan)):
        cent[name] = O.voxel_grid(O.crop(scan, BOX)[0], 0.2)[0]
        assert np.array_equal(got[f"{name}_filtered"][:, :3].view(np.uint32), cent[name].view(np.uint32))
    merged_xyz = np.concatenate([O.transform(cent["robot"], *inp.robot_tf), O.transform(cent["zx120"], *inp.zx120_tf)])
    assert np.array_equal(xyz(got["merged"]).view(np.uint32), merged_xyz.view(np.uint32))
    # generator on the merged cloud (colours: robot red, backhoe blue, a = 255)
    colours = np.ascontiguousarray(got["merged"][:, 16:20]).view(np.uint32).ravel()
    assert set(np.unique(colours)) == {0xFFFF0000, 0xFF0000FF}
    prm = A.excavation_params()
    ns, na = A.excavation_counts(prm)
    ref_t, ref_a, ri = O.excavate(merged_xyz, colours, np.zeros(3), np.array([0.0, 0.0, 0.0, 1.0]), prm, ns, na)
    info = got["excavation_info"]
    assert (info.n_kept, info.n_surface, info.n_area) == (ri.n_kept, ri.n_surface, ri.n_area)
    assert 0 < merged_xyz.shape[0] - info.n_kept          # the scans saw the pit: something was carved
    dz = assert_clouds_match(got["terrain"], ref_t, int(ri.n_kept), "terrain") + assert_clouds_match(got["area"], ref_a, 0, "area")
    # virtual_lidar setup on the generator's own outputs
    area_xyz, terrain_xyz = xyz(got["area"]), xyz(got["terrain"])
    onorm = O.compute_normals(area_xyz, 1.5)
    ok = np.isfinite(onorm).all(1) & np.isfinite(got["normals"]).all(1)
    assert ok.mean() > 0.99
    assert np.percentile(np.abs(np.sum(got["normals"][ok].astype(np.float64) * onorm[ok], 1)), 1) > 0.99999
    ocells, ob = O.generate_cells(area_xyz, got["normals"], 0.1, 10, 1.5)
    cells = got["cells"]
    assert cells.size == ocells.size > 1000
    for k in ("x", "y", "z"):
        assert np.array_equal(cells[k], ocells[k])
    for k in ("nx", "ny", "nz"):
        assert np.max(np.abs(cells[k] - ocells[k])) <= 1.2e-7
    tix = O.Index(terrain_xyz, O.GRID)
    ocands = O.generate_candidates(tix, ob, got["fixed"], 64, 3.0, 1.1)
    cands = got["candidates"]
    assert cands.size == ocands.size > 20
    for k in ("x", "y", "z"):
        assert np.array_equal(cands[k], ocands[k])
    # sweep on the device's cells and candidates, observed cloud = the zx120 filter output
    ref = O.sweep(tix, O.Index(cent["zx120"], O.GRID), cells, A.default_params(), got["fixed"], cands, want_mask=True)   # one thread: the stale-flag rule is sequential
    r = got["sweep"]
    assert np.array_equal(r.mask, ref["mask"])
    assert_same_optimum(r.best_index, ref)
    assert np.allclose(r.scores, ref["scores"], rtol=1e-12, atol=0)
    assert r.counters["samples"] == ref["counters"]["samples"]
    assert np.array_equal(r.cell_flags, ref["flags"])
    # calc_drivable_area on the robot's filtered scan
    rxy = scenes.ROBOT_BASE["xyz"][:2]
    ogrid = O.drivable_area(cent["robot"], inp.robot_tf[0], inp.robot_tf[1], rxy, rxy)
    assert np.array_equal(got["occupancy_grid"], ogrid) and (ogrid == -1).any()
    assert all(v > 0 for v in got["device_ms"].values()), got["device_ms"]
    print({k: round(v, 3) for k, v in got["device_ms"].items()}, "z last-bit differences:", dz)


def test_launch_graph_full_size(ctx):
    """Config 4's scan size (2 x 1 M points): the graph runs end to end and its products keep their invariants."""
    inp = scenes.pipeline_inputs(n_scan=1_000_000, sweep="tiny")
    got = run_launch_graph(ctx, inp, num_candidates=32 * 32, grid_resolution=0.05, vertical_layers=20)
    info = got["excavation_info"]
    prm = A.excavation_params()
    ns, na = A.excavation_counts(prm)
    assert got["terrain"].shape[0] == info.n_kept + ns and got["area"].shape[0] == na
    assert got["cells"].size > 10_000 and got["candidates"].size > 500
    assert got["sweep"].best_index >= 0 and np.isfinite(got["sweep"].best_score)
    n = np.stack([got["cells"]["nx"], got["cells"]["ny"], got["cells"]["nz"]], 1).astype(np.float64)
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-6)
    print({k: round(v, 3) for k, v in got["device_ms"].items()})


@pytest.mark.parametrize("n_scan,kw", [(150_000, dict(num_candidates=64)),
                                       (1_000_000, dict(num_candidates=256, grid_resolution=0.05, vertical_layers=12))])
def test_resident_graph_equals_host_buffer_graph(ctx, n_scan, kw):
    """with every intermediate cloud left in device buffers (read in place by the next stage) the graph
    produces the same bytes as with host buffers between the stages."""
    inp = scenes.pipeline_inputs(n_scan=n_scan, sweep="tiny")
    host = run_launch_graph(ctx, inp, mask=True, **kw)
    dev = run_launch_graph_resident(ctx, inp, download=True, mask=True, **kw)
    for k in ("robot_filtered", "zx120_filtered", "merged", "terrain", "area"):
        assert np.array_equal(np.ascontiguousarray(host[k]).view(np.uint8).reshape(-1), dev[k].view(np.uint8).reshape(-1)), k
    assert np.array_equal(host["normals"].view(np.uint32), dev["normals"].view(np.uint32))
    assert np.array_equal(host["cells"].view(np.uint8), dev["cells"].view(np.uint8))
    assert np.array_equal(host["candidates"].view(np.uint8), dev["candidates"].view(np.uint8))
    assert np.array_equal(host["sweep"].mask, dev["sweep"].mask) and host["sweep"].best_index == dev["sweep"].best_index
    assert np.array_equal(host["sweep"].scores, dev["sweep"].scores)
    assert np.array_equal(host["occupancy_grid"], dev["occupancy_grid"])
    print(n_scan, "resident graph wall ms", round(dev["wall_ms"], 2), dev["counts"])


@pytest.mark.parametrize("resident", [False, True])
def test_cpp_host_runs_the_same_graph(ctx, tmp_path, resident):
    """examples/launch_graph.cpp drives the C ABI from C++ (the reference's language) with no Python in between: same
    scans in, same sizes, optimal pose and score out as nodes.run_launch_graph."""
    import json
    import subprocess
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    exe = root / "examples" / "_build" / "launch_graph"
    if not exe.exists():
        subprocess.check_call(["make", "-C", str(root / "examples")])
    inp = scenes.pipeline_inputs(n_scan=150_000, sweep="tiny")
    inp.robot_scan.astype("<f4").tofile(tmp_path / "robot.f32")
    inp.zx120_scan.astype("<f4").tofile(tmp_path / "zx120.f32")
    cmd = [str(exe), "--robot", str(tmp_path / "robot.f32"), "--zx120", str(tmp_path / "zx120.f32"), "--candidates", "64"]
    out = subprocess.run(cmd + (["--resident"] if resident else []), capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    got = json.loads(out.stdout.strip().splitlines()[-1])
    ref = run_launch_graph(ctx, inp, num_candidates=64)
    assert got["resident"] is resident
    assert got["robot_filtered"] == ref["robot_filtered"].shape[0] and got["zx120_filtered"] == ref["zx120_filtered"].shape[0]
    assert got["merged"] == ref["merged"].shape[0] and got["terrain"] == ref["terrain"].shape[0] and got["area"] == ref["area"].shape[0]
    assert got["kept"] == ref["excavation_info"].n_kept
    assert got["cells"] == ref["cells"].size and got["candidates"] == ref["candidates"].size
    r = ref["sweep"]
    assert got["best_index"] == r.best_index and got["rays_visible"] == r.counters["rays_visible"]
    assert got["best_score"] == pytest.approx(r.best_score, rel=1e-12)
    b = ref["candidates"][r.best_index]
    assert got["best_pose"] == [b["x"], b["y"], b["z"]]
    assert got["grid_occupied"] == int((ref["occupancy_grid"] == 100).sum()) and got["grid_unknown"] == int((ref["occupancy_grid"] == -1).sum())
    assert got["kernels_launched"] > 50
