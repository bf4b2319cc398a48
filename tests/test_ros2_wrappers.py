"""The rclcpp wrappers in ros2/src, compiled and driven without ROS2.

There is no ROS2 in this image, so the five node executables a maintainer drops into the reference package are built
against tests/mock_ros2 (rclcpp / tf2_ros / message types reduced to what the wrappers use, an in-process bus and a
script player behind rclcpp::spin) and linked with libplc.so.  The CPU test checks that they compile and link; the GPU
tests play message / TF / timer / parameter scripts through them and compare what they publish with the same calls made
through the ctypes driver: topics, headers, wire layouts, the reference's skip rules (no TF, empty cloud, disabled),
parameter re-reads."""
import json
import os
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
MOCK = ROOT / "tests" / "mock_ros2"
BUILD = ROOT / "tests" / "_build"
NODES = ["virtual_lidar", "pointcloud_filter", "pointcloud_merger", "excavated_surface_generator", "calc_drivable_area"]
GEO = ROOT / "oracle" / "ref_shim" / "GeographicLib"       # stand-in LocalCartesian.hpp: the only header pointcloud_merger needs beyond the mock


def build(name: str, stub: str | None = None) -> Path:
    """stub: a tests/mock_ros2/plc_stub_*.cpp whose definitions shadow libplc.so's (host-logic tests on a box without a GPU)."""
    exe = BUILD / (f"ros2_{name}" + ("_stub" if stub else ""))
    srcs = [ROOT / "ros2" / "src" / f"{name}_node.cpp", MOCK / "mock_ros2.cpp"] + ([MOCK / stub] if stub else [])
    deps = srcs + [MOCK / "mock_ros2.hpp", ROOT / "ros2" / "src" / "cloud_msg.hpp", ROOT / "include" / "plc.h",
                   ROOT / "include" / "plc_node.h"]
    if exe.exists() and all(exe.stat().st_mtime >= d.stat().st_mtime for d in deps):
        return exe
    BUILD.mkdir(exist_ok=True)
    lib = ROOT / "pointcloud_processor_b200" / "lib"
    geo = []
    if name == "pointcloud_merger":                        # <GeographicLib/LocalCartesian.hpp> without the rest of oracle/ref_shim
        (BUILD / "geo").mkdir(exist_ok=True)
        if not (BUILD / "geo" / "GeographicLib").exists():
            (BUILD / "geo" / "GeographicLib").symlink_to(GEO)
        geo = [f"-I{BUILD / 'geo'}"]
    cmd = ["g++", "-O1", "-std=c++17", "-Wall", "-Wextra", "-Werror", f"-I{MOCK}", f"-I{ROOT / 'include'}",
           f"-I{ROOT / 'ros2' / 'src'}", *geo, *map(str, srcs), "-o", str(exe), f"-L{lib}", "-lplc", f"-Wl,-rpath,{lib}"]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return exe


@pytest.mark.parametrize("name", NODES)
def test_wrapper_compiles_against_mock_ros2(name):
    try:
        exe = build(name)
    except subprocess.CalledProcessError as e:          # show the compiler's words
        pytest.fail(e.stderr[-4000:])
    assert exe.exists()


def play(name: str, script: list[str], tmp: Path, stub: str | None = None):
    """Runs one wrapper executable over a script; returns (published messages as (meta, bytes), log lines)."""
    out = tmp / f"out_{name}_{len(list(tmp.glob('out_*')))}"
    out.mkdir()
    (out / "script.txt").write_text("\n".join(script) + "\n")
    env = dict(os.environ, MOCK_ROS_SCRIPT=str(out / "script.txt"), MOCK_ROS_OUT=str(out))
    r = subprocess.run([str(build(name, stub))], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    msgs = []
    for j in sorted(out.glob("[0-9][0-9][0-9].json")):
        b = j.with_suffix(".bin")
        msgs.append((json.loads(j.read_text()), b.read_bytes() if b.exists() else b""))
    log = (out / "log.txt").read_text().splitlines() if (out / "log.txt").exists() else []
    return msgs, log


def cloud_line(tmp: Path, topic: str, rec: np.ndarray, frame: str, stamp=(12, 34), fields="x:0 y:4 z:8") -> str:
    rec = np.ascontiguousarray(rec)
    f = tmp / f"cloud_{len(list(tmp.glob('cloud_*')))}.bin"
    rec.tofile(f)
    n = rec.shape[0]
    step = rec.nbytes // n if n else rec.shape[1] * rec.itemsize
    return f"cloud {topic} {f} {n} {step} {frame} {stamp[0]} {stamp[1]} {fields}"


def quat_zyx(yaw, pitch):
    cy, sy, cp, sp = np.cos(yaw / 2), np.sin(yaw / 2), np.cos(pitch / 2), np.sin(pitch / 2)
    return np.array([-sy * sp, cy * sp, sy * cp, cy * cp])


def tf_line(target, source, t, q):
    return f"tf {target} {source} " + " ".join(repr(float(v)) for v in (*t, *q))


PC2_XYZ = [["x", 0, 7, 1], ["y", 4, 7, 1], ["z", 8, 7, 1]]


@pytest.mark.gpu
def test_pointcloud_filter_node(ctx, tmp_path):
    from pointcloud_processor_b200.nodes import filter_scan
    rng = np.random.default_rng(5)
    robot = np.zeros((120_000, 4), np.float32)                      # x y z intensity, the velodyne driver's layout
    robot[:, :3] = rng.uniform(-20, 20, (robot.shape[0], 3))
    robot[::97, 1] = np.nan
    back = np.zeros((80_000, 6), np.float32)                        # x, y, z NOT at 0/4/8: the fields are looked up by name
    back[:, 1:4] = rng.uniform(-18, 18, (back.shape[0], 3))
    msgs, log = play("pointcloud_filter", [
        "param voxel_leaf_size double 0.1", "param robot_front_range double 12.0", "param backhoe_height_range double 6.0",
        cloud_line(tmp_path, "/four_wheel_robot/velodyne_points", robot, "four_wheel_robot/velodyne", (100, 7)),
        cloud_line(tmp_path, "/zx120/velodyne_points", back, "zx120/velodyne", (101, 8), "x:4 y:8 z:12"),
        cloud_line(tmp_path, "/zx120/velodyne_points", np.zeros((0, 4), np.float32), "zx120/velodyne", (102, 9)),
    ], tmp_path)
    assert [m["topic"] for m, _ in msgs] == ["/four_wheel_robot/filtered_points", "/zx120/filtered_points", "/zx120/filtered_points"]
    want = [filter_scan(ctx, robot, 12.0, 10.0, 10.0, 0.1), filter_scan(ctx, back, 15.0, 10.0, 6.0, 0.1, 24, (4, 8, 12))]
    for (m, b), w, frame, stamp in zip(msgs, want, ("four_wheel_robot/velodyne", "zx120/velodyne"), ([100, 7], [101, 8])):
        assert (m["frame_id"], m["stamp"]) == (frame, stamp)        # header copied from the input (:79)
        assert (m["point_step"], m["height"], m["width"], m["fields"]) == (16, 1, w.shape[0], PC2_XYZ) and w.shape[0] > 100
        assert np.array_equal(np.frombuffer(b, np.uint32), w.view(np.uint32).ravel())
    m, b = msgs[2]                                                  # an empty scan is published as an empty cloud (Q7)
    assert (m["width"], m["bytes"], m["stamp"]) == (0, 0, [102, 9]) and b == b""
    assert "Simplified Scan Matcher initialized" in " ".join(log)


@pytest.mark.gpu
def test_calc_drivable_area_node(ctx, tmp_path):
    from pointcloud_processor_b200 import _abi as A
    rng = np.random.default_rng(6)
    n = 150_000
    pts = np.zeros((n, 4), np.float32)
    pts[:, :2] = rng.uniform(-25, 25, (n, 2))
    pts[:, 2] = 0.2 * np.sin(0.3 * pts[:, 0]) - 2.0 + np.where(rng.random(n) < 0.15, rng.uniform(0, 2.5, n), 0.0)
    t, q = np.array([9.25, -3.0, 2.0]), quat_zyx(2.6, 0.1245)
    scan = cloud_line(tmp_path, "/four_wheel_robot/filtered_points", pts, "four_wheel_robot/velodyne")
    empty = cloud_line(tmp_path, "/four_wheel_robot/filtered_points", np.zeros((0, 4), np.float32), "four_wheel_robot/velodyne")
    msgs, log = play("calc_drivable_area", [
        "param grid_resolution double 0.5", "param map_width double 60.0", "param map_height double 50.0",
        "param min_points_per_cell int 4",
        scan,                                                                        # no TF yet: skipped with a warning
        tf_line("map", "four_wheel_robot/velodyne", t, q), tf_line("map", "four_wheel_robot/base_link", (9.0, -3.0, 0.0), (0, 0, 0, 1)),
        scan, empty,
        tf_line("map", "four_wheel_robot/base_link", (11.0, -2.0, 0.0), (0, 0, 0, 1)),    # the robot moved: the start stays
        scan,
    ], tmp_path)
    assert [m["type"] for m, _ in msgs] == ["OccupancyGrid", "OccupancyGrid"]
    prm = A.DrivableParams()
    A.lib().plc_drivable_params_default(prm)
    prm.grid_resolution, prm.map_width, prm.map_height, prm.min_points_per_cell = 0.5, 60.0, 50.0, 4
    for (m, b), robot in zip(msgs, ((9.0, -3.0), (11.0, -2.0))):
        want, info = ctx.drivable_area(pts, t, q, robot, (9.0, -3.0), prm, 16)
        assert (m["width"], m["height"], m["frame_id"], m["origin_w"]) == (120, 100, "map", 1.0)
        assert m["origin"] == [robot[0] - 30.0, robot[1] - 25.0, 0.0] and m["resolution"] == 0.5
        assert np.array_equal(np.frombuffer(b, np.int8).reshape(100, 120), want)
    text = "\n".join(log)
    assert "Transform from four_wheel_robot/velodyne to map not available yet" in text
    assert "Received empty point cloud" in text and text.count("Start position set at (9.00, -3.00)") == 1


@pytest.mark.gpu
def test_excavated_surface_generator_node(ctx, tmp_path):
    from pointcloud_processor_b200 import _abi as A, scenes
    xyz = scenes.terrain("L", 40_000, seed=3, side=24.0)
    rec = np.zeros((xyz.shape[0], 8), np.float32)                   # PointXYZRGB: rgb (bit-cast) at byte 16
    rec[:, :3] = xyz
    rec[:, 4] = np.full(xyz.shape[0], 0xFFFF0000, np.uint32).view(np.float32)
    line = cloud_line(tmp_path, "/matched_point_cloud", rec, "map", (55, 66), "x:0 y:4 z:8 rgb:16")
    t, q = (0.5, -0.25, 0.1), quat_zyx(0.3, 0.0)
    # disabled: the input is republished untouched (:264-267)
    msgs, _ = play("excavated_surface_generator", ["param excavation.enabled bool false", line], tmp_path)
    assert len(msgs) == 1 and msgs[0][0]["topic"] == "/excavated_terrain" and msgs[0][1] == rec.tobytes()
    assert msgs[0][0]["stamp"] == [55, 66]
    # enabled: no TF -> republished with a warning (:274-279); TF -> carved terrain + /excavation_area; a parameter
    # changed while running is picked up by the 1 s timer (:69-71)
    msgs, log = play("excavated_surface_generator", [
        "param excavation.depth double 2.5", line, tf_line("map", "zx120/base_link", t, q), line,
        "setparam excavation.depth double 1.5", line, "timers", line], tmp_path)
    assert [m["topic"] for m, _ in msgs] == ["/excavated_terrain"] + ["/excavated_terrain", "/excavation_area"] * 3
    assert msgs[0][1] == rec.tobytes() and "Could not get zx120 transform" in "\n".join(log)
    prm = A.excavation_params()
    for depth, k in ((2.5, 1), (2.5, 3), (1.5, 5)):                 # the change shows only after the timer has fired
        prm.depth = depth
        terrain, area, info = ctx.excavate(rec, t, q, prm)
        (mt, bt), (ma, ba) = msgs[k], msgs[k + 1]
        assert (mt["width"], mt["point_step"], mt["frame_id"], mt["stamp"]) == (terrain.shape[0], 32, "map", [55, 66])
        assert (ma["width"], ma["point_step"], ma["frame_id"], ma["stamp"]) == (area.shape[0], 32, "map", [55, 66])
        assert mt["fields"] == PC2_XYZ + [["rgb", 16, 7, 1]]
        assert bt == terrain.tobytes() and ba == area.tobytes()
        assert info.n_kept < rec.shape[0] and info.n_surface > 0


@pytest.mark.gpu
def test_virtual_lidar_node(ctx, tmp_path):
    from pointcloud_processor_b200 import scenes
    from pointcloud_processor_b200.nodes import VirtualLidarNode

    def records(xyz):
        r = np.zeros((xyz.shape[0], 8), np.float32)
        r[:, :3] = xyz
        return r
    terrain = records(scenes.terrain("rect", 30_000, seed=21, side=24.0))
    area = records(scenes.excavation_area("rect", 0.05)[0])
    base = (0.25, -0.5, 0.0)
    msgs, log = play("virtual_lidar", [
        "param num_candidates int 49", "param vertical_layers int 6",
        cloud_line(tmp_path, "/excavation_area", area, "map", fields="x:0 y:4 z:8 rgb:16"),      # -> grid markers (:286)
        cloud_line(tmp_path, "/excavated_terrain", terrain, "map", fields="x:0 y:4 z:8 rgb:16"),
        "timers",                                                                                # TF lookup fails: nothing
        tf_line("map", "zx120/base_link", base, (0, 0, 0, 1)),
        "timers",
        "setparam num_candidates int 25", "timers",                                              # re-read every cycle
    ], tmp_path)
    node = VirtualLidarNode(ctx)
    node.set_parameter("num_candidates", 49)
    node.set_parameter("vertical_layers", 6)
    node.update_parameters()
    node.excavation_area_callback(area)
    node.terrain_callback(terrain)
    cells0, flags0 = node.cells()
    node.set_base_transform(base)
    res = [node.run_optimization()]
    cands = [node.candidates()]
    grids = [node.cells()]
    node.set_parameter("num_candidates", 25)
    res.append(node.run_optimization())
    cands.append(node.candidates())
    grids.append(node.cells())
    assert res[0].ran == 1 and res[1].ran == 1 and cands[0].size != cands[1].size

    topics = [m["topic"] for m, _ in msgs]
    cycle = ["/optimal_mobile_lidar_position", "/mobile_lidar_candidate_positions", "/excavation_grid_visualization"]
    assert topics == ["/excavation_grid_visualization"] + cycle * 2

    def rows(b):
        return np.frombuffer(b, np.float64).reshape(-1, 14)
    g0 = rows(msgs[0][1])                                           # DELETEALL + one cube per cell, before any sweep
    assert g0.shape[0] == cells0.size + 1 and g0[0, 0] == 3 and np.all(g0[1:, 1] == 1)
    assert np.array_equal(g0[1:, 3], cells0["x"]) and np.array_equal(g0[1:, 5], cells0["z"])
    assert np.allclose(g0[1:, 6:9], 0.1 * 0.6) and msgs[0][0]["frame_id"] == "map"
    for k in range(2):
        (mb, _), (mc, bc), (mg, bg) = msgs[1 + 3 * k: 4 + 3 * k]
        r = res[k]
        assert mb["point"] == [r.best.x, r.best.y, r.best.z] and mb["frame_id"] == "map"
        c = rows(bc)                                                # DELETEALL, zx120 cube, candidates, optimum cylinder
        assert c.shape[0] == cands[k].size + 3 and mc["namespaces"] == ["", "zx120_lidar", "mobile_lidar_candidates", "optimal_mobile_lidar"]
        assert np.array_equal(c[2:-1, 3], cands[k]["x"]) and np.array_equal(c[2:-1, 4], cands[k]["y"]) and np.array_equal(c[2:-1, 5], cands[k]["z"])
        assert list(c[1, 3:6]) == [r.fixed.x, r.fixed.y, r.fixed.z] and list(c[-1, 3:6]) == [r.best.x, r.best.y, r.best.z]
        assert c[-1, 1] == 3 and np.all(c[2:-1, 1] == 2) and np.all(c[2:-1, 13] == 8.0)
        g = rows(bg)[1:]                                            # colours follow the six flags (:881-925)
        cells, flags = grids[k]
        rng_ok, fov_ok, vis_ok = (flags & 9) > 0, (flags & 18) > 0, (flags & 36) > 0
        want = np.where(~rng_ok[:, None], [0, 0, 1.0], np.where(~fov_ok[:, None], [1.0, 1.0, 0], np.where(~vis_ok[:, None], [1.0, 0, 0], [0, 1.0, 0])))
        assert g.shape[0] == cells.size and np.array_equal(g[:, 9:12], want)
        assert (want[:, 1] == 1).any() and (want[:, 0] == 1).any()
    text = "\n".join(log)
    assert "sweep failed" not in text and "Failed to process" not in text
