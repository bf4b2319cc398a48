"""Golden vectors from the reference's OWN source: runs oracle/_ref/<node>_ref — /root/reference/src/{virtual_lidar,
pointcloud_filter, pointcloud_merger, calc_drivable_area, excavated_surface_generator}.cpp, unmodified, compiled against stand-in PCL / tf2 /
ROS2 headers (`make -C oracle ref`) — on seeded inputs and stores what the nodes publish and log in
tests/golden/refsrc_<name>.npz, next to the inputs (or their SHA-256 where tests/refsrc_inputs.py rebuilds them).  Needs /root/reference (this container);
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


def play(node, make_lines):
    """Runs oracle/_ref/<node>_ref over the script `make_lines(tmp)` returns; (published messages, log text)."""
    subprocess.check_call(["make", "-C", os.fspath(ROOT / "oracle"), "ref"], stdout=subprocess.DEVNULL)
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        (tmp / "script.txt").write_text("\n".join(make_lines(tmp)) + "\n")
        env = dict(os.environ, MOCK_ROS_SCRIPT=os.fspath(tmp / "script.txt"), MOCK_ROS_OUT=os.fspath(tmp))
        subprocess.run([os.fspath(ROOT / "oracle" / "_ref" / f"{node}_ref")], env=env, check=True, capture_output=True)
        msgs = []
        for j in sorted(tmp.glob("[0-9][0-9][0-9].json")):
            b = j.with_suffix(".bin")
            msgs.append((json.loads(j.read_text()), b.read_bytes() if b.exists() else b""))
        log = (tmp / "log.txt").read_text() if (tmp / "log.txt").exists() else ""
    return msgs, log


def cloud_line(tmp, topic, rec, frame, stamp, fields):
    f = tmp / f"cloud_{len(list(tmp.glob('cloud_*')))}.bin"
    np.ascontiguousarray(rec).tofile(f)
    step = rec.shape[1] * rec.itemsize
    return f"cloud {topic} {f} {rec.shape[0]} {step} {frame} {stamp[0]} {stamp[1]} {fields}"


def tf_line(target, source, t, q):
    return f"tf {target} {source} " + " ".join(repr(float(v)) for v in (*t, *q))


def run_reference(area, terrain, observed, num_candidates, layers, base, res=None):
    def lines(tmp):
        out = [f"param num_candidates int {num_candidates}", f"param vertical_layers int {layers}"]
        if res is not None:
            out.append(f"param grid_resolution double {res!r}")
        for topic, cloud in (("/excavation_area", area), ("/excavated_terrain", terrain), ("/zx120/filtered_points", observed)):
            if cloud is not None:
                out.append(cloud_line(tmp, topic, cloud, "map", (10, 20), "x:0 y:4 z:8 rgb:16"))
        return out + [tf_line("map", "zx120/base_link", base, (0, 0, 0, 1)), "timers"]
    return play("virtual_lidar", lines)


def param_lines(params):
    return [f"param {k} {'int' if isinstance(v, int) else 'double'} {v!r}" for k, v in params.items()]


def other_nodes():
    """pointcloud_filter, calc_drivable_area and excavated_surface_generator of the reference on the seeded inputs of
    tests/refsrc_inputs.py."""
    sys.path.insert(0, os.fspath(ROOT / "tests"))
    import refsrc_inputs as R
    gold = ROOT / "tests" / "golden"
    # ---- pointcloud_filter (src/pointcloud_filter.cpp): crop + VoxelGrid of both scans
    scans = R.filter_scans()
    msgs, _ = play("pointcloud_filter", lambda tmp: param_lines(R.FILTER_PARAMS) + [
        cloud_line(tmp, "/four_wheel_robot/velodyne_points", scans["robot"], "four_wheel_robot/velodyne", (100, 7), "x:0 y:4 z:8 intensity:12"),
        cloud_line(tmp, "/zx120/velodyne_points", scans["backhoe"], "zx120/velodyne", (101, 8), "x:0 y:4 z:8 intensity:12")])
    assert [m["topic"] for m, _ in msgs] == ["/four_wheel_robot/filtered_points", "/zx120/filtered_points"]
    out = {}
    for who, (m, b) in zip(("robot", "backhoe"), msgs):
        out[f"{who}_records"] = np.frombuffer(b, np.float32).reshape(-1, m["point_step"] // 4).copy()
        out[f"{who}_meta"] = json.dumps(m)
        out[f"{who}_input_sha256"] = R.digest(scans[who])
    np.savez_compressed(gold / "refsrc_filter.npz", **out)
    print("filter", {k: v.shape for k, v in out.items() if hasattr(v, "shape") and v.ndim == 2})
    # ---- pointcloud_merger (src/pointcloud_merger.cpp) on what the filter node just published
    filt = {who: out[f"{who}_records"] for who in ("robot", "backhoe")}
    msgs, log = play("pointcloud_merger", lambda tmp: R.merger_script(tmp, filt, cloud_line, tf_line))
    topics = [m["topic"] for m, _ in msgs]
    assert topics == R.MERGER_TOPICS, topics
    np.savez_compressed(gold / "refsrc_merger.npz", topics=json.dumps(topics), metas=json.dumps([m for m, _ in msgs]), log=log,
                        robot_only=np.frombuffer(msgs[0][1], np.uint8).reshape(-1, 32).copy(), merged=np.frombuffer(msgs[4][1], np.uint8).reshape(-1, 32).copy(),
                        robot_colored=np.frombuffer(msgs[5][1], np.uint8).reshape(-1, 32).copy(), backhoe_colored=np.frombuffer(msgs[6][1], np.uint8).reshape(-1, 32).copy(),
                        same_again=np.array([msgs[2][1] == msgs[0][1], msgs[3][1] == msgs[1][1], msgs[1][1] == msgs[0][1]]),
                        robot_input_sha256=R.digest(filt["robot"]), backhoe_input_sha256=R.digest(filt["backhoe"]))
    print("merger", topics, msgs[4][0]["width"], [m for m, _ in msgs[7:]])
    # ---- calc_drivable_area (src/calc_drivable_area.cpp)
    scan, (t, q), robot = R.drivable_scan()
    line = lambda tmp: cloud_line(tmp, "/four_wheel_robot/filtered_points", scan, "four_wheel_robot/velodyne", (5, 6), "x:0 y:4 z:8 intensity:12")
    msgs, log = play("calc_drivable_area", lambda tmp: param_lines(R.DRIVABLE_PARAMS) + [
        line(tmp),                                                           # no TF yet: nothing published
        tf_line("map", "four_wheel_robot/velodyne", t, q), tf_line("map", "four_wheel_robot/base_link", (*robot, 0.0), (0, 0, 0, 1)),
        line(tmp),
        tf_line("map", "four_wheel_robot/base_link", (robot[0] + 2.0, robot[1] + 1.0, 0.0), (0, 0, 0, 1)), line(tmp)])
    assert [m["type"] for m, _ in msgs] == ["OccupancyGrid", "OccupancyGrid"]
    np.savez_compressed(gold / "refsrc_drivable.npz", grid0=np.frombuffer(msgs[0][1], np.int8).copy(), grid1=np.frombuffer(msgs[1][1], np.int8).copy(),
                        meta0=json.dumps(msgs[0][0]), meta1=json.dumps(msgs[1][0]), log=log, input_sha256=R.digest(scan))
    print("drivable", msgs[0][0]["width"], msgs[0][0]["height"], {v: int((np.frombuffer(msgs[0][1], np.int8) == v).sum()) for v in (-1, 0, 100)})
    # ---- excavated_surface_generator (src/excavated_surface_generator.cpp)
    for shape in ("L", "rect"):
        rec, (bt, bq) = R.generator_cloud(shape)
        msgs, log = play("excavated_surface_generator", lambda tmp: [f"param excavation.l_shape_enabled bool {'true' if shape == 'L' else 'false'}",
                                                                     tf_line("map", "zx120/base_link", bt, bq),
                                                                     cloud_line(tmp, "/matched_point_cloud", rec, "map", (55, 66), "x:0 y:4 z:8 rgb:16")])
        by = {m["topic"]: (m, b) for m, b in msgs}
        tm, tb = by["/excavated_terrain"]
        am, ab = by["/excavation_area"]
        np.savez_compressed(gold / f"refsrc_generator_{shape}.npz", terrain=np.frombuffer(tb, np.uint8).reshape(-1, tm["point_step"]).copy(),
                            area=np.frombuffer(ab, np.uint8).reshape(-1, am["point_step"]).copy(), terrain_meta=json.dumps(tm),
                            area_meta=json.dumps(am), input_sha256=R.digest(rec), topics=json.dumps([m["topic"] for m, _ in msgs]))
        print("generator", shape, "terrain", tm["width"], "area", am["width"], [m["topic"] for m, _ in msgs])


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


def parse_cycles(msgs, log):
    """One dict per optimisation cycle of a script with several `timers` lines: a cycle publishes optimum, candidates
    and grid, in that order, and logs both statistics blocks once (an area message publishes a grid of its own)."""
    rows = lambda b: np.frombuffer(b, np.float64).reshape(-1, 14)
    topics = [m["topic"] for m, _ in msgs]
    at = [i for i, t in enumerate(topics) if t == "/optimal_mobile_lidar_position"]
    assert all(topics[i + 1] == "/mobile_lidar_candidate_positions" and topics[i + 2] == "/excavation_grid_visualization" for i in at), topics
    blocks = log.split("INFO ZX120 LiDAR Only Evaluation")[1:]
    assert len(blocks) == len(at)
    out = []
    for i, text in zip(at, blocks):
        grid, cand = rows(msgs[i + 2][1])[1:], rows(msgs[i + 1][1])
        num = lambda pat: [float(x) for x in re.findall(pat, text)]
        counts = {c: [int(v) for v in re.findall(rf"{c} \([^)]*\): (\d+) cells", text)] for c in ("Green", "Red", "Blue", "Yellow")}
        out.append(dict(cells_xyz=grid[:, 3:6].copy(), cell_rgb=grid[:, 9:12].copy(), cands_xyz=cand[2:-1, 3:6].copy(),
                        fixed_xyz=cand[1, 3:6].copy(), best_xyz=np.array(msgs[i][0]["point"]),
                        fixed_total_2dp=num(r"Total Score \(ZX120 only\): ([-\d.]+)")[-1], best_score_2dp=num(r"\nINFO Total Score: ([-\d.]+)")[-1],
                        colour_counts=np.array([[counts[c][i] for c in ("Green", "Red", "Blue", "Yellow")] for i in (0, 1)])))
    return out


def cycles_case():
    """tests/golden/refsrc_cycles.npz: five optimisation cycles of ONE node process — what carries over between cycles
    (the per-cell flags, the cells until the next area message, parameters as of the last cycle's re-read)."""
    sys.path.insert(0, os.fspath(ROOT / "tests"))
    import refsrc_inputs as R
    C = R.cycles_inputs(scenes)
    msgs, log = play("virtual_lidar", lambda tmp: R.cycles_script(tmp, C, cloud_line, tf_line))
    cyc = parse_cycles(msgs, log)
    flat = {f"c{k}_{name}": v for k, c in enumerate(cyc) for name, v in c.items()}
    path = ROOT / "tests" / "golden" / "refsrc_cycles.npz"
    np.savez_compressed(path, n_cycles=len(cyc), log=log, **{f"{k}_sha256": R.digest(v) for k, v in C.items()}, **flat)
    for k, c in enumerate(cyc):
        print("cycle", k, "cells", c["cells_xyz"].shape[0], "candidates", c["cands_xyz"].shape[0], "best", c["best_xyz"], "score", c["best_score_2dp"],
              "fixed", c["fixed_total_2dp"], "colours", c["colour_counts"].tolist())
    print("->", path.relative_to(ROOT), f"{path.stat().st_size / 1e3:.0f} kB")


def vl_edges_case():
    """tests/golden/refsrc_vl_edges.npz: virtual_lidar's guards — empty terrain / observed / area messages, the tree that
    outlives its cloud, a missing transform — as six cycles of one process (and two timer ticks that publish nothing)."""
    sys.path.insert(0, os.fspath(ROOT / "tests"))
    import refsrc_inputs as R
    V = R.vl_edges_inputs(scenes)
    msgs, log = play("virtual_lidar", lambda tmp: R.vl_edges_script(tmp, V, cloud_line, tf_line))
    cyc = parse_cycles(msgs, log)
    assert len(cyc) == 6, len(cyc)
    flat = {f"c{k}_{name}": v for k, c in enumerate(cyc) for name, v in c.items()}
    path = ROOT / "tests" / "golden" / "refsrc_vl_edges.npz"
    np.savez_compressed(path, n_cycles=len(cyc), log=log, topics=json.dumps([m["topic"] for m, _ in msgs]),
                        **{f"{k}_sha256": R.digest(v) for k, v in V.items()}, **flat)
    for k, c in enumerate(cyc):
        print("vl edge cycle", k, "cells", c["cells_xyz"].shape[0], "candidates", c["cands_xyz"].shape[0], "cand z", np.unique(c["cands_xyz"][:, 2])[:3],
              "score", c["best_score_2dp"], "fixed", c["fixed_total_2dp"], "colours", c["colour_counts"].tolist())
    print("->", path.relative_to(ROOT), f"{path.stat().st_size / 1e3:.0f} kB")


def variants_case():
    """tests/golden/refsrc_variants.npz: the nodes around the sweep with non-default parameters (tests/refsrc_inputs.py
    VARIANT_*): filter ranges and leaf, drivable grid resolution / gradient / clear radius, both pit shapes with other
    dimensions, slopes, densities, offsets and a rotated base."""
    sys.path.insert(0, os.fspath(ROOT / "tests"))
    import refsrc_inputs as R
    out = {}
    names = ["pointcloud_filter", "calc_drivable_area", "excavated_surface_generator_L", "excavated_surface_generator_rect"]
    for name in names:
        node = name if not name.startswith("excavated") else "excavated_surface_generator"
        msgs, log = play(node, lambda tmp: R.variants_scripts(tmp, cloud_line, tf_line)[name])
        for m, b in msgs:
            if m["topic"] == "/excavation_markers":
                continue
            key = f"{name}{m['topic'].replace('/', '__')}"
            assert key not in out, key
            out[key] = np.frombuffer(b, np.uint8).copy()
            out[key + "__meta"] = json.dumps(m)
        print(name, [(m["topic"], m.get("width")) for m, _ in msgs])
    path = ROOT / "tests" / "golden" / "refsrc_variants.npz"
    np.savez_compressed(path, **out)
    print("->", path.relative_to(ROOT), f"{path.stat().st_size / 1e3:.0f} kB")


def config_case(name, num_candidates=None):
    """tests/golden/refsrc_<name>.npz: a BASELINE.json configuration through the reference's own virtual_lidar.cpp — the
    scene of scenes.CONFIGS[name] with the node's own lattices (every valid cell, where bench scenes truncate to exactly
    V x H).  cfg1 (configs[0], the reference's own CPU-runnable case: L-shaped pit, 50 k points, 16 x 16 requested poses,
    16 x 900 rays per pose) runs in full; cfg2 (configs[1], the configuration the metric is quoted on: rectangular pit,
    200 k points, 32 x 1800 rays per pose) with 36 requested poses of its first height layer."""
    cfg = scenes.CONFIGS[name]
    layers = scenes.LAYERS_FOUND[name]
    num = num_candidates or cfg["px"] * cfg["px"]
    terrain = records(scenes.terrain(cfg["shape"], cfg["n"], cfg["seed"], cfg["side"]))
    area = records(scenes.excavation_area(cfg["shape"], 0.05)[0])
    out = parse(*run_reference(area, terrain, None, num, layers, (0.0, 0.0, 0.0), res=cfg["res"]))
    out.pop("log")
    path = ROOT / "tests" / "golden" / f"refsrc_{name}.npz"
    sys.path.insert(0, os.fspath(ROOT / "tests"))
    import refsrc_inputs as R
    np.savez_compressed(path, num_candidates=num, vertical_layers=layers, grid_resolution=cfg["res"],
                        terrain_sha256=R.digest(terrain), area_sha256=R.digest(area),
                        cell_rgb_code=(out.pop("cell_rgb") @ np.array([4, 2, 1])).astype(np.uint8), cells_xyz=out.pop("cells_xyz").astype(np.float64), **out)
    print(name, "cells", np.load(path)["cells_xyz"].shape[0], "candidates", out["cands_xyz"].shape[0], "best", out["best_xyz"], "score", out["best_score_2dp"],
          "fixed", out["fixed_total_2dp"], "colours", out["colour_counts"].tolist(), "->", path.relative_to(ROOT), f"{path.stat().st_size / 1e3:.0f} kB")


def edges_case():
    """tests/golden/refsrc_edges.npz: the skip / republish / empty-cloud / parameter-timing rules of the nodes around the
    sweep, as the reference's own sources apply them to the scripts of tests/refsrc_inputs.edges_scripts: per script the
    published messages' metadata, a SHA-256 of every payload, and the log."""
    import hashlib
    sys.path.insert(0, os.fspath(ROOT / "tests"))
    import refsrc_inputs as R
    E = R.edges_inputs(scenes)
    out = {f"{k}_sha256": R.digest(v) for k, v in E.items()}
    for name in ("pointcloud_filter", "calc_drivable_area", "excavated_surface_generator_disabled", "excavated_surface_generator"):
        msgs, log = play(name.replace("_disabled", ""), lambda tmp: R.edges_scripts(tmp, E, cloud_line, tf_line)[name])
        out[name] = json.dumps(dict(metas=[m for m, _ in msgs], payload_sha256=[hashlib.sha256(b).hexdigest() for _, b in msgs], log=log.splitlines()))
        print(name, [(m["topic"], m.get("width")) for m, _ in msgs])
    path = ROOT / "tests" / "golden" / "refsrc_edges.npz"
    np.savez_compressed(path, **out)
    print("->", path.relative_to(ROOT), f"{path.stat().st_size / 1e3:.0f} kB")


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
    cycles_case()
    vl_edges_case()
    config_case("cfg1")
    config_case("cfg2", 36)
    variants_case()
    edges_case()
    other_nodes()
