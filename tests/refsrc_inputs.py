"""Seeded inputs of the reference-source fixtures for the nodes around the sweep (tests/golden/refsrc_*.npz): built
here, identically, by tools/make_refsrc_golden.py (which feeds them to the reference's own nodes) and by the tests
(which feed them to the oracle and to the CUDA path).  The fixtures store a SHA-256 of each input next to the outputs."""
import hashlib

import numpy as np


def digest(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def quat_zyx(yaw, pitch):
    cy, sy, cp, sp = np.cos(yaw / 2), np.sin(yaw / 2), np.cos(pitch / 2), np.sin(pitch / 2)
    return np.array([-sy * sp, cy * sp, sy * cp, cy * cp])


def filter_scans():
    """Two LiDAR scans in their sensor frames: x y z intensity records (point_step 16), some non-finite points, a few
    points exactly on the crop box's faces."""
    out = {}
    for name, seed, n in (("robot", 31, 40_000), ("backhoe", 32, 30_000)):
        rng = np.random.Generator(np.random.PCG64(seed))
        p = np.zeros((n, 4), np.float32)
        p[:, :3] = rng.uniform(-22, 22, (n, 3))
        p[:, 2] = rng.uniform(-3, 12, n)
        p[:, 3] = rng.uniform(0, 255, n)
        p[rng.choice(n, 25, replace=False), rng.integers(0, 3, 25)] = np.nan
        p[:6, :3] = [[0.0, 1.0, 1.0], [15.0, 1.0, 1.0], [3.0, 10.0, 1.0], [3.0, -10.0, 1.0], [3.0, 1.0, -1.5], [3.0, 1.0, 10.0]]
        out[name] = p
    return out


FILTER_PARAMS = {"voxel_leaf_size": 0.2, "robot_front_range": 12.0, "backhoe_height_range": 6.0}


# pointcloud_merger: its inputs are the filter fixture's outputs; map <- */velodyne_link as the scripted TF tree answers
MERGER_NAMES = {"robot": "four_wheel_robot", "backhoe": "zx120"}
MERGER_TF = {"robot": (np.array([9.25, -3.0, 2.0]), quat_zyx(2.6, 0.1245)), "backhoe": (np.array([0.55, 0.4, 3.5]), quat_zyx(0.0, 0.4363))}
MERGER_RGB = {"robot": (255, 0, 0), "backhoe": (0, 0, 255)}
MERGER_TOPICS = ["/matched_point_cloud", "/robot_colored_point_cloud"] * 2 + ["/matched_point_cloud", "/robot_colored_point_cloud",
                                                                             "/backhoe_colored_point_cloud", "/tf", "/tf", "/tf"]


def merger_script(tmp, clouds, cloud_line, tf_line):
    """The message / TF / timer script both the reference's pointcloud_merger.cpp and ros2/src/pointcloud_merger_node.cpp
    are played: ticks before the GNSS origin, with one source, with a source whose TF is missing, with both; then the
    GNSS half (fix + heading -> map -> base_link broadcasts for both vehicles)."""
    cl = lambda who: cloud_line(tmp, f"/{MERGER_NAMES[who]}/filtered_points", clouds[who], f"{MERGER_NAMES[who]}/velodyne", (100, 7), "x:0 y:4 z:8")
    tfl = lambda who: tf_line("map", f"{MERGER_NAMES[who]}/velodyne_link", *MERGER_TF[who])
    return [cl("robot"), tfl("robot"), "timers",                                 # no GNSS origin yet (:309): nothing
            "fix /zx120/gnss_compass_front/fix 35.0 139.0 40.0 0", "timers",     # origin set; robot cloud only
            cl("backhoe"), "timers",                                             # no TF for the backhoe's cloud: robot only + a warning (:389-393)
            tfl("backhoe"), "timers",                                            # both
            "heading /four_wheel_robot/gnss_compass/calc_heading 0 0 0.6 0.8",
            tf_line("four_wheel_robot/gnss_link", "four_wheel_robot/base_link", (-0.5, 0, 0), (0, 0, 0, 1)),
            "fix /four_wheel_robot/gnss_compass_front/fix 35.00005 139.00008 41.0 0",      # map -> four_wheel_robot/base_link, no backhoe heading
            tf_line("zx120/base_link", "zx120/gnss_link", (2.0, 0, 3.8), (0, 0, 0, 1)),
            "heading /zx120/gnss_compass/calc_heading 0 0 0.1 0.99498743710662",           # map -> zx120/base_link: identity
            "fix /four_wheel_robot/gnss_compass_front/fix 35.00006 139.00007 41.5 0"]      # relative heading + antenna offset


def drivable_scan():
    """A filtered robot scan in the sensor frame (point_step 16) and the two transforms the node looks up."""
    rng = np.random.Generator(np.random.PCG64(33))
    n = 60_000
    p = np.zeros((n, 4), np.float32)
    p[:, :2] = rng.uniform(-25, 25, (n, 2))
    p[:, 2] = 0.2 * np.sin(0.3 * p[:, 0]) * np.cos(0.2 * p[:, 1]) - 2.0 + np.where(rng.random(n) < 0.15, rng.uniform(0, 2.5, n), 0.0)
    p[rng.choice(n, 20, replace=False), rng.integers(0, 3, 20)] = np.nan
    return p, (np.array([9.25, -3.0, 2.0]), quat_zyx(2.6, 0.1245)), (9.0, -3.0)


DRIVABLE_PARAMS = {"grid_resolution": 0.5, "map_width": 60.0, "map_height": 50.0, "min_points_per_cell": 4}


def generator_cloud(shape):
    """/matched_point_cloud: un-carved rolling ground around where the pit will be, PointXYZRGB records (point_step 32)."""
    rng = np.random.Generator(np.random.PCG64(34 if shape == "L" else 35))
    n = 9_000
    r = np.zeros((n, 8), np.float32)
    r[:, :2] = rng.uniform(-2.0, 7.0, (n, 2))
    r[:, 2] = 0.15 * np.sin(0.4 * r[:, 0]) * np.cos(0.3 * r[:, 1]) + rng.normal(0, 0.01, n) - rng.uniform(0, 1.2, n) * (rng.random(n) < 0.3)
    r[:, 4] = np.where(rng.random(n) < 0.5, np.uint32(0xFFFF0000), np.uint32(0xFF0000FF)).astype(np.uint32).view(np.float32)
    base = (np.array([0.3, -0.2, 0.0]), np.array([0.0, 0.0, np.sin(0.2), np.cos(0.2)]) if shape == "rect" else np.array([0.0, 0.0, 0.0, 1.0]))
    return r, base


# ---- virtual_lidar over several cycles of one process (tests/golden/refsrc_cycles.npz) -------------------------------
CYCLES_BASES = [(0.0, 0.0, 0.1), (0.3, 0.2, 0.05)]
# parameters in force DURING each cycle's sweep (re-read at the top of runOptimization, :457) and the ones the cell
# lattice was generated with (the area callback uses the values of the LAST re-read, :172 -> :236-287)
CYCLES_PARAMS = [dict(num_candidates=36, vertical_layers=6, grid_resolution=0.1, sensor_height=1.1, search_radius=3.0, max_distance=15.0),
                 dict(num_candidates=25, vertical_layers=6, grid_resolution=0.1, sensor_height=1.6, search_radius=2.0, max_distance=4.5),
                 dict(num_candidates=25, vertical_layers=6, grid_resolution=0.1, sensor_height=1.6, search_radius=2.0, max_distance=4.5),
                 dict(num_candidates=25, vertical_layers=4, grid_resolution=0.2, sensor_height=1.6, search_radius=2.0, max_distance=9.0),
                 dict(num_candidates=25, vertical_layers=4, grid_resolution=0.2, sensor_height=1.6, search_radius=2.0, max_distance=9.0)]
CYCLES_LATTICE = [(0.1, 6), (0.1, 6), (0.1, 6), (0.1, 6), (0.2, 4)]       # (grid_resolution, vertical_layers) behind each cycle's cells
CYCLES_FRESH_CELLS = [True, False, False, True, True]                      # an area message arrived before this cycle: flags start from zero


def cycles_inputs(scenes):
    def rec(xyz):
        r = np.zeros((xyz.shape[0], 8), np.float32)
        r[:, :3] = xyz[:, :3]
        return r
    area = rec(scenes.excavation_area("rect", 0.05)[0])
    return dict(area=area, terrain_a=rec(scenes.terrain("rect", 30_000, seed=7, side=24.0)),
                terrain_b=rec(scenes.terrain("rect", 24_000, seed=8, side=24.0)), observed=rec(area[::7, :3].copy()))


def cycles_script(tmp, C, cloud_line, tf_line):
    cl = lambda topic, key: cloud_line(tmp, topic, C[key], "map", (10, 20), "x:0 y:4 z:8 rgb:16")
    sp = lambda name, kind, v: f"setparam {name} {kind} {v!r}"
    return ["param num_candidates int 36", "param vertical_layers int 6",
            cl("/excavation_area", "area"), cl("/excavated_terrain", "terrain_a"), tf_line("map", "zx120/base_link", CYCLES_BASES[0], (0, 0, 0, 1)),
            "timers",                                                                                    # cycle 0
            sp("max_distance", "double", 4.5), sp("sensor_height", "double", 1.6), sp("search_radius", "double", 2.0), sp("num_candidates", "int", 25),
            "timers",                                                                                    # cycle 1: new parameters, same data, flags carried over
            cl("/excavated_terrain", "terrain_b"), cl("/zx120/filtered_points", "observed"), tf_line("map", "zx120/base_link", CYCLES_BASES[1], (0, 0, 0, 1)),
            "timers",                                                                                    # cycle 2: new terrain, an observed cloud, the base moved
            sp("grid_resolution", "double", 0.2), sp("vertical_layers", "int", 4), sp("max_distance", "double", 9.0),
            cl("/excavation_area", "area"),                                                              # lattice still at 0.1 x 6: the re-read comes with the cycle
            "timers",                                                                                    # cycle 3
            cl("/excavation_area", "area"),                                                              # now 0.2 x 4
            "timers"]                                                                                    # cycle 4


# ---- edge rules of the nodes around the sweep (tests/golden/refsrc_edges.npz): the scripts tests/test_ros2_wrappers.py
# plays through the drop-in executables, played through the reference's own nodes -----------------------------------
def edges_inputs(scenes):
    rng = np.random.Generator(np.random.PCG64(5))
    robot = np.zeros((60_000, 4), np.float32)                       # x y z intensity, the velodyne driver's layout
    robot[:, :3] = rng.uniform(-20, 20, (robot.shape[0], 3))
    robot[::97, 1] = np.nan
    back = np.zeros((40_000, 6), np.float32)                        # x, y, z NOT at 0/4/8: the fields are looked up by name
    back[:, 1:4] = rng.uniform(-18, 18, (back.shape[0], 3))
    far = np.zeros((3_000, 4), np.float32)                          # nothing inside the crop box: an empty cloud goes out
    far[:, :3] = rng.uniform(-20, -1, (far.shape[0], 3))
    rng = np.random.Generator(np.random.PCG64(6))
    n = 80_000
    pts = np.zeros((n, 4), np.float32)
    pts[:, :2] = rng.uniform(-25, 25, (n, 2))
    pts[:, 2] = 0.2 * np.sin(0.3 * pts[:, 0]) - 2.0 + np.where(rng.random(n) < 0.15, rng.uniform(0, 2.5, n), 0.0)
    xyz = scenes.terrain("L", 20_000, seed=3, side=24.0)
    rec = np.zeros((xyz.shape[0], 8), np.float32)                   # PointXYZRGB: rgb (bit-cast) at byte 16
    rec[:, :3] = xyz
    rec[:, 4] = np.full(xyz.shape[0], 0xFFFF0000, np.uint32).view(np.float32)
    return dict(robot=robot, back=back, far=far, pts=pts, rec=rec)


EDGES_DRIVABLE_TF = (np.array([9.25, -3.0, 2.0]), quat_zyx(2.6, 0.1245))
EDGES_GENERATOR_TF = ((0.5, -0.25, 0.1), quat_zyx(0.3, 0.0))
EDGES_FILTER_PARAMS = {"voxel_leaf_size": 0.1, "robot_front_range": 12.0, "backhoe_height_range": 6.0}
EDGES_DRIVABLE_PARAMS = {"grid_resolution": 0.5, "map_width": 60.0, "map_height": 50.0, "min_points_per_cell": 4}


def edges_scripts(tmp, E, cloud_line, tf_line):
    empty = np.zeros((0, 4), np.float32)
    scan = lambda a=None: cloud_line(tmp, "/four_wheel_robot/filtered_points", E["pts"] if a is None else a, "four_wheel_robot/velodyne", (12, 34), "x:0 y:4 z:8")
    line = cloud_line(tmp, "/matched_point_cloud", E["rec"], "map", (55, 66), "x:0 y:4 z:8 rgb:16")
    t, q = EDGES_DRIVABLE_TF
    gt, gq = EDGES_GENERATOR_TF
    prm = lambda d: [f"param {k} {'int' if isinstance(v, int) else 'double'} {v!r}" for k, v in d.items()]
    return {
        "pointcloud_filter": prm(EDGES_FILTER_PARAMS) + [
            cloud_line(tmp, "/four_wheel_robot/velodyne_points", E["robot"], "four_wheel_robot/velodyne", (100, 7), "x:0 y:4 z:8"),
            cloud_line(tmp, "/zx120/velodyne_points", E["back"], "zx120/velodyne", (101, 8), "x:4 y:8 z:12"),
            cloud_line(tmp, "/zx120/velodyne_points", empty, "zx120/velodyne", (102, 9), "x:0 y:4 z:8"),      # an empty scan
            cloud_line(tmp, "/four_wheel_robot/velodyne_points", E["far"], "four_wheel_robot/velodyne", (103, 10), "x:0 y:4 z:8")],
        "calc_drivable_area": prm(EDGES_DRIVABLE_PARAMS) + [
            scan(),                                                                      # no TF yet: skipped with a warning
            tf_line("map", "four_wheel_robot/velodyne", t, q), tf_line("map", "four_wheel_robot/base_link", (9.0, -3.0, 0.0), (0, 0, 0, 1)),
            scan(), scan(empty),
            tf_line("map", "four_wheel_robot/base_link", (11.0, -2.0, 0.0), (0, 0, 0, 1)),    # the robot moved: the start stays
            scan()],
        "excavated_surface_generator_disabled": ["param excavation.enabled bool false", line],
        "excavated_surface_generator": ["param excavation.depth double 2.5", line, tf_line("map", "zx120/base_link", gt, gq), line,
                                        "setparam excavation.depth double 1.5", line, "timers", line],
    }


# ---- virtual_lidar's guards (tests/golden/refsrc_vl_edges.npz): empty messages, stale trees, a missing transform -----
def vl_edges_inputs(scenes):
    def rec(xyz):
        r = np.zeros((xyz.shape[0], 8), np.float32)
        r[:, :3] = xyz[:, :3]
        return r
    area = rec(scenes.excavation_area("rect", 0.05)[0])
    return dict(area=area, terrain=rec(scenes.terrain("rect", 30_000, seed=23, side=24.0)), observed=rec(area[::7, :3].copy()))


def vl_edges_script(tmp, V, cloud_line, tf_line):
    empty = np.zeros((0, 8), np.float32)
    cl = lambda topic, a: cloud_line(tmp, topic, a, "map", (10, 20), "x:0 y:4 z:8 rgb:16")
    return ["param num_candidates int 49", "param vertical_layers int 5",
            cl("/excavation_area", V["area"]), tf_line("map", "zx120/base_link", (0.0, 0.0, 0.0), (0, 0, 0, 1)),
            "timers",                                              # no terrain message at all: the guard holds (:455), nothing
            cl("/excavated_terrain", empty), "timers",             # cycle 0: a cloud but no tree: every gated ray visible, ground height 0
            cl("/excavated_terrain", V["terrain"]), "timers",      # cycle 1
            cl("/excavated_terrain", empty), "timers",             # cycle 2: the tree stays, ground heights go to 0
            cl("/zx120/filtered_points", V["observed"]), "timers", # cycle 3: the observed-cloud branch of the fixed sensor
            cl("/zx120/filtered_points", empty), "timers",         # cycle 4: an empty observed cloud: back to the ray path
            cl("/excavation_area", empty), "timers",               # cycle 5: an empty area message leaves the grid in place (:168)
            "notf map zx120/base_link", "timers"]                  # no transform: nothing


# ---- non-default parameters of the nodes around the sweep (tests/golden/refsrc_variants.npz) -------------------------
VARIANT_GENERATORS = {
    "L": (dict(depth=0.7, slope_angle=60.0, offset_x=3.5, offset_y=0.5, point_density=0.08, terrain_search_radius=0.6, l_shape_enabled=True,
               arm1_length=2.4, arm1_width=1.0, arm2_length=1.6, arm2_width=1.4),
          ((0.2, 0.3, 0.05), (0.0, 0.0, float(np.sin(0.2)), float(np.cos(0.2))))),
    "rect": (dict(depth=1.3, slope_angle=82.0, offset_x=4.5, offset_y=-0.5, point_density=0.06, l_shape_enabled=False, width=1.5, length=2.2),
             ((-0.3, 0.1, 0.0), (0.0, 0.0, float(np.sin(-0.15)), float(np.cos(-0.15))))),
}
VARIANT_DRIVABLE = dict(grid_resolution=0.25, map_width=40.0, map_height=30.0, max_gradient=0.5, min_points_per_cell=2, start_clear_radius=5.0)
VARIANT_FILTER = dict(voxel_leaf_size=0.05, robot_front_range=8.0, robot_side_range=6.0, robot_height_range=4.0,
                      backhoe_front_range=10.0, backhoe_side_range=7.0, backhoe_height_range=3.0)


def variant_cloud(name):
    """Rolling ground over the place the pit of VARIANT_GENERATORS[name] will be carved, PointXYZRGB records."""
    rng = np.random.Generator(np.random.PCG64(41 if name == "L" else 42))
    n = 7_000
    r = np.zeros((n, 8), np.float32)
    r[:, :2] = rng.uniform(-1.0, 8.0, (n, 2)) + np.array([0.0, -3.5])
    r[:, 2] = 0.15 * np.sin(0.4 * r[:, 0]) * np.cos(0.3 * r[:, 1]) + rng.normal(0, 0.01, n) - rng.uniform(0, 1.5, n) * (rng.random(n) < 0.3)
    r[:, 4] = np.where(rng.random(n) < 0.5, np.uint32(0xFFFF0000), np.uint32(0xFF0000FF)).astype(np.uint32).view(np.float32)
    return r


def variants_scripts(tmp, cloud_line, tf_line):
    prm = lambda d, pre="": [f"param {pre}{k} {'bool' if isinstance(v, bool) else 'int' if isinstance(v, int) else 'double'} "
                             f"{str(v).lower() if isinstance(v, bool) else repr(v)}" for k, v in d.items()]
    scans = filter_scans()
    scan, (t, q), robot = drivable_scan()
    out = {
        "pointcloud_filter": prm(VARIANT_FILTER) + [
            cloud_line(tmp, "/four_wheel_robot/velodyne_points", scans["robot"], "four_wheel_robot/velodyne", (100, 7), "x:0 y:4 z:8 intensity:12"),
            cloud_line(tmp, "/zx120/velodyne_points", scans["backhoe"], "zx120/velodyne", (101, 8), "x:0 y:4 z:8 intensity:12")],
        "calc_drivable_area": prm(VARIANT_DRIVABLE) + [
            tf_line("map", "four_wheel_robot/velodyne", t, q), tf_line("map", "four_wheel_robot/base_link", (*robot, 0.0), (0, 0, 0, 1)),
            cloud_line(tmp, "/four_wheel_robot/filtered_points", scan, "four_wheel_robot/velodyne", (5, 6), "x:0 y:4 z:8 intensity:12")],
    }
    for name, (p, (bt, bq)) in VARIANT_GENERATORS.items():
        out[f"excavated_surface_generator_{name}"] = prm(p, "excavation.") + [
            tf_line("map", "zx120/base_link", bt, bq), cloud_line(tmp, "/matched_point_cloud", variant_cloud(name), "map", (55, 66), "x:0 y:4 z:8 rgb:16")]
    return out
