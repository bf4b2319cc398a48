"""Synthetic inputs for the virtual_lidar sweep, shaped like the reference's own data products
(SURVEY §8 D2).  Everything is deterministic given the seed; the same arrays feed the CUDA path
and the CPU oracle.

  /excavated_terrain  ground height field on a jittered lattice with the pit carved out, plus the
                      pit surface (bottom + slopes) laid out with the rules of
                      excavated_surface_generator.cpp:487-584 (generateExcavatedSurface)
  /excavation_area    target pit surface per excavated_surface_generator.cpp:350-455
                      (generateExcavationArea), with analytic per-point normals
  cells               reference lattice rule (virtual_lidar.cpp:236-340) via plc_host_generate_cells
  candidates          reference lattice rule (virtual_lidar.cpp:550-625) via plc_host_generate_candidates,
                      repeated for P_z sensor heights {1.1 + 0.4k}

The reference's generator cannot run at these sizes (it rebuilds a KD-tree per point),
so the ground height is analytic: z = 0.15 sin(0.4x) cos(0.3y) (+-1 cm noise on ground points).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from . import _abi as A
from .context import host_generate_candidates, host_generate_cells

# excavated_surface_generator.cpp:28-47 defaults
DEPTH = 1.0
SLOPE_DEG = 75.0
PIT_CENTER = (4.0, 1.0)          # offset_x, offset_y from zx120/base_link at the map origin
FIXED_OFFSET = (0.4, 0.5, 3.5)   # virtual_lidar.cpp:103-105
FIXED_PITCH = -np.pi / 6         # :106


def ground_height(x, y):
    return 0.15 * np.sin(0.4 * x) * np.cos(0.3 * y)


def pit_boxes(shape: str):
    """getExcavationBoxes (excavated_surface_generator.cpp:138-181): (min_x, max_x, min_y, max_y) per box."""
    if shape == "L":
        a1l, a1w, a2l, a2w = 2.0, 1.2, 2.0, 1.2
        arm1 = (-a1w / 2, a1w / 2, -a1l / 2 - a1l / 2, -a1l / 2 + a1l / 2)
        c2x, c2y = a2l / 2, -a1l + a2w / 2
        arm2 = (c2x - a2l / 2, c2x + a2l / 2, c2y - a2w / 2, c2y + a2w / 2)
        return [arm1, arm2]
    if shape == "rect":
        length, width = 1.8, 1.2
        return [(-length / 2, length / 2, -width / 2, width / 2)]
    raise ValueError(shape)


def _inside_any(x, y, boxes):
    m = np.zeros(np.broadcast(x, y).shape, bool)
    for (x0, x1, y0, y1) in boxes:
        m |= (x >= x0) & (x <= x1) & (y >= y0) & (y <= y1)
    return m


def _edge_lattice(boxes, d):
    """Lattice of generateExcavationArea/generateExcavatedSurface: points inside any box, the outer-edge
    subset (isOuterEdge with tolerance = point_density) and each edge point's outward direction."""
    x0 = min(b[0] for b in boxes); x1 = max(b[1] for b in boxes)
    y0 = min(b[2] for b in boxes); y1 = max(b[3] for b in boxes)
    nx = int((x1 - x0) / d) + 1
    ny = int((y1 - y0) / d) + 1
    xs = x0 + np.arange(nx + 1) * d
    ys = y0 + np.arange(ny + 1) * d
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    X = X.ravel(); Y = Y.ravel()
    inside = _inside_any(X, Y, boxes)
    X = X[inside]; Y = Y[inside]
    out_px = ~_inside_any(X + d, Y, boxes); out_mx = ~_inside_any(X - d, Y, boxes)
    out_py = ~_inside_any(X, Y + d, boxes); out_my = ~_inside_any(X, Y - d, boxes)
    edge = out_px | out_mx | out_py | out_my
    dir_x = np.where(out_px, 1.0, np.where(out_mx, -1.0, 0.0))
    dir_y = np.where(out_py, 1.0, np.where(out_my, -1.0, 0.0))
    return X, Y, edge, dir_x, dir_y


def excavation_area(shape: str, density: float = 0.05):
    """/excavation_area cloud (yaw 0) and analytic unit normals (n_z >= 0) per point."""
    boxes = pit_boxes(shape)
    slope_offset = DEPTH / np.tan(np.deg2rad(SLOPE_DEG))
    X, Y, edge, dx, dy = _edge_lattice(boxes, density)
    cx, cy = PIT_CENTER
    gx, gy = cx + X, cy + Y
    h = ground_height(gx, gy)
    pts = [np.stack([gx, gy, h - DEPTH], 1)]
    nrm = [np.tile([0.0, 0.0, 1.0], (gx.size, 1))]
    n_depth = int(DEPTH / density)
    ex, ey, eh, edx, edy = X[edge], Y[edge], h[edge], dx[edge], dy[edge]
    for k in range(1, n_depth):
        off = slope_offset * (k / n_depth)
        pts.append(np.stack([cx + ex + edx * off, cy + ey + edy * off, eh - DEPTH + k * density], 1))
        # slope face rises outward: normal points inward and up
        n = np.stack([-edx * DEPTH, -edy * DEPTH, np.full(ex.size, slope_offset)], 1)
        n /= np.linalg.norm(n, axis=1, keepdims=True)
        nrm.append(n)
    return np.concatenate(pts).astype(np.float32), np.concatenate(nrm).astype(np.float32)


def pit_surface(shape: str, density: float):
    """generateExcavatedSurface: bottom lattice + slope lines from every outer-edge lattice point."""
    boxes = pit_boxes(shape)
    slope_offset = DEPTH / np.tan(np.deg2rad(SLOPE_DEG))
    X, Y, edge, dx, dy = _edge_lattice(boxes, density)
    cx, cy = PIT_CENTER
    gx, gy = cx + X, cy + Y
    pts = [np.stack([gx, gy, ground_height(gx, gy) - DEPTH], 1)]
    n_slope = int(slope_offset / density) + 1
    ex, ey, edx, edy = X[edge], Y[edge], dx[edge], dy[edge]
    for k in range(n_slope + 1):
        zr = k / n_slope
        sx, sy = cx + ex + edx * slope_offset * zr, cy + ey + edy * slope_offset * zr
        pts.append(np.stack([sx, sy, ground_height(sx, sy) - DEPTH * (1.0 - zr)], 1))
    return np.concatenate(pts)


def _pit_count(shape, density):
    return pit_surface(shape, density).shape[0]


def terrain(shape: str, n_points: int, seed: int, side: float = 30.0, pit_fraction: float = 0.10):
    """/excavated_terrain with exactly n_points points (float32 xyz)."""
    rng = np.random.Generator(np.random.PCG64(seed))      # an explicit bit generator: the stream is fixed by NumPy's compatibility policy
    # pit density so that the pit surface is ~pit_fraction of the cloud
    target = max(64, int(n_points * pit_fraction))
    lo, hi = 0.002, 0.5
    for _ in range(40):
        mid = np.sqrt(lo * hi)
        if _pit_count(shape, mid) > target:
            lo = mid
        else:
            hi = mid
    pit = pit_surface(shape, hi)
    n_ground = n_points - pit.shape[0]
    if n_ground < 0:
        pit = pit[rng.permutation(pit.shape[0])[:n_points]]
        n_ground = 0
    boxes = pit_boxes(shape)
    slope_offset = DEPTH / np.tan(np.deg2rad(SLOPE_DEG))
    cx, cy = PIT_CENTER
    area = side * side - sum((b[1] - b[0] + 2 * slope_offset) * (b[3] - b[2] + 2 * slope_offset) for b in boxes)
    pitch = np.sqrt(area / max(n_ground, 1)) * 0.985     # a few percent spare, trimmed below
    m = int(np.ceil(side / pitch))
    ii, jj = np.meshgrid(np.arange(m), np.arange(m), indexing="ij")
    gx = cx - side / 2 + (ii.ravel() + 0.5 + rng.uniform(-0.3, 0.3, ii.size)) * pitch
    gy = cy - side / 2 + (jj.ravel() + 0.5 + rng.uniform(-0.3, 0.3, jj.size)) * pitch
    # carve: drop ground inside the pit's top outline (boxes grown by the slope offset,
    # isInsideExcavationArea at z_relative = 0, excavated_surface_generator.cpp:327-348)
    keep = np.ones(gx.size, bool)
    for (x0, x1, y0, y1) in boxes:
        keep &= ~((gx - cx >= x0 - slope_offset) & (gx - cx <= x1 + slope_offset) &
                  (gy - cy >= y0 - slope_offset) & (gy - cy <= y1 + slope_offset))
    gx, gy = gx[keep], gy[keep]
    if gx.size < n_ground:
        raise RuntimeError(f"ground lattice too sparse: {gx.size} < {n_ground}")
    sel = np.sort(rng.permutation(gx.size)[:n_ground])
    gx, gy = gx[sel], gy[sel]
    gz = ground_height(gx, gy) + rng.uniform(-0.01, 0.01, gx.size)
    cloud = np.concatenate([np.stack([gx, gy, gz], 1), pit]).astype(np.float32)
    assert cloud.shape[0] == n_points
    return np.ascontiguousarray(cloud)


@dataclass
class Scene:
    name: str
    terrain: np.ndarray        # (N,3) float32
    area: np.ndarray           # (A,3) float32
    area_normals: np.ndarray   # (A,3) float32
    cells: np.ndarray          # (M,) CELL_DTYPE
    bounds: A.GridBounds
    fixed: np.ndarray          # (1,) POSE_DTYPE
    candidates: np.ndarray     # (C,) POSE_DTYPE
    params: A.Params
    meta: dict


def fixed_pose() -> np.ndarray:
    f = np.zeros(1, A.POSE_DTYPE)
    f["x"], f["y"], f["z"] = FIXED_OFFSET          # zx120/base_link at the map origin
    f["pitch"], f["yaw"] = FIXED_PITCH, 0.0
    return f


def make_cells(area, normals, m_target: int | None, res: float, layers: int | None, host_generate_cells=host_generate_cells):
    if layers is None:
        # smallest vertical_layers whose lattice yields at least m_target valid cells
        lo, hi = 1, 8
        while host_generate_cells(area, None, res, hi, 1.5)[0].size < m_target:
            lo, hi = hi, hi * 2
            if hi > 4096:
                raise RuntimeError("cannot reach the requested cell count")
        while lo < hi:
            mid = (lo + hi) // 2
            if host_generate_cells(area, None, res, mid, 1.5)[0].size >= m_target:
                hi = mid
            else:
                lo = mid + 1
        layers = hi
    cells, bounds = host_generate_cells(area, normals, res, layers, 1.5)
    if m_target is not None:
        if cells.size < m_target:
            raise RuntimeError(f"lattice res={res} layers={layers} gives {cells.size} < {m_target} cells")
        cells = cells[:m_target].copy()    # truncate in reference order (D1)
    make_cells.last_layers = layers
    return cells, bounds


def make_candidates(terrain_xyz, bounds, fixed, px: int, pz: int, search_radius=3.0, base_height=1.1,
                    host_generate_candidates=host_generate_candidates):
    layers = [host_generate_candidates(terrain_xyz, bounds, fixed, px * px, search_radius, base_height + 0.4 * k)
              for k in range(pz)]
    return np.concatenate(layers) if layers else np.zeros(0, A.POSE_DTYPE)


# BASELINE.json configs -> concrete sizes.  (res, layers) chosen so that the reference lattice
# yields at least M valid cells; the list is then truncated to exactly M = V*H.
CONFIGS = {
    # name: shape, N, px, pz, M, seed, side, res, layers
    "tiny": dict(shape="L", n=6_000, px=4, pz=1, m=None, seed=11, side=14.0, res=0.2, layers=4),
    "small": dict(shape="rect", n=20_000, px=6, pz=1, m=None, seed=12, side=16.0, res=0.1, layers=6),
    "ref_default": dict(shape="L", n=50_000, px=10, pz=1, m=None, seed=13, side=30.0, res=0.1, layers=10),
    # layers=None: smallest vertical_layers giving >= m valid cells (then truncated to exactly m)
    "cfg1": dict(shape="L", n=50_000, px=16, pz=1, m=16 * 900, seed=1, side=30.0, res=0.05, layers=None),
    "cfg2": dict(shape="rect", n=200_000, px=32, pz=2, m=32 * 1800, seed=2, side=30.0, res=0.025, layers=None),
    "cfg3": dict(shape="L", n=1_000_000, px=64, pz=4, m=32 * 1800, seed=3, side=30.0, res=0.03, layers=None),
    "cfg5": dict(shape="L", n=4_000_000, px=128, pz=8, m=64 * 2048, seed=6, side=40.0, res=0.025, layers=None),
}
# what the search for the smallest sufficient vertical_layers finds (checked by tests/test_oracle.py); a caller whose
# cell generator is slow (the single-threaded oracle in bench.py's reference arm) passes these instead of searching
LAYERS_FOUND = {"cfg1": 18, "cfg2": 74, "cfg3": 50, "cfg5": 99}


def build(name: str, setup=None, **override) -> Scene:
    """`setup` = (generate_cells, generate_candidates) replaces the plc_host_* steps with functions of the same
    signature (bench.py's reference arm passes the oracle's, so that it never loads the product library)."""
    cfg = dict(CONFIGS[name])
    cfg.update(override)
    gen_cells, gen_cands = setup if setup is not None else (host_generate_cells, host_generate_candidates)
    terr = terrain(cfg["shape"], cfg["n"], cfg["seed"], cfg["side"])
    area, normals = excavation_area(cfg["shape"], 0.05)
    fixed = fixed_pose()
    cells, bounds = make_cells(area, normals, cfg["m"], cfg["res"], cfg["layers"], gen_cells)
    cands = make_candidates(terr, bounds, fixed, cfg["px"], cfg["pz"], host_generate_candidates=gen_cands)
    prm = A.default_params()
    meta = dict(cfg)
    meta.update(layers_used=int(round((bounds.max_z - bounds.min_z) / max(cells["z"][1] - cells["z"][0], 1e-12))) if cells.size > 1 and cells["x"][0] == cells["x"][1] else None,
                layers_found=int(make_cells.last_layers), n_cells=int(cells.size), n_candidates=int(cands.size), n_terrain=int(terr.shape[0]),
                n_area=int(area.shape[0]))
    return Scene(name, terr, area, normals, cells, bounds, fixed, cands, prm, meta)


# ---- config 4: two LiDAR scans -> pointcloud_filter -> pointcloud_merger -> virtual_lidar ------------------
# Static transforms base_link -> velodyne_link, "x y z yaw pitch roll" (launch/tf_zx120.launch.xml:4,
# launch/tf_four_wheel_robot.launch.xml:4), and the map poses of the two bases: zx120/base_link is pinned to
# the map origin (pointcloud_merger.cpp:218-228); the robot stands beside the pit, facing it.
ZX120_VELODYNE = dict(xyz=(0.55, 0.4, 3.5), yaw=0.0, pitch=0.4363)
ROBOT_VELODYNE = dict(xyz=(0.25, 0.0, 2.0), yaw=0.0, pitch=0.1245)
ROBOT_BASE = dict(xyz=(9.0, -3.0, 0.0), yaw=2.6)


def _quat_yp(yaw: float, pitch: float) -> np.ndarray:
    """Quaternion (x, y, z, w) of Rz(yaw) * Ry(pitch), the composition static_transform_publisher applies."""
    cy, sy, cp, sp = np.cos(yaw / 2), np.sin(yaw / 2), np.cos(pitch / 2), np.sin(pitch / 2)
    return np.array([-sy * sp, cy * sp, sy * cp, cy * cp])


def _rot(q: np.ndarray) -> np.ndarray:
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def sensor_in_map(base: dict, mount: dict):
    """lookupTransform("map", "<name>/velodyne_link") (pointcloud_merger.cpp:362-366): translation, quaternion."""
    qb = _quat_yp(base.get("yaw", 0.0), 0.0)
    qm = _quat_yp(mount["yaw"], mount["pitch"])
    t = np.asarray(base["xyz"], float) + _rot(qb) @ np.asarray(mount["xyz"], float)
    x1, y1, z1, w1 = qb
    x2, y2, z2, w2 = qm
    q = np.array([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                  w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2])
    return t, q


def lidar_scan(t: np.ndarray, q: np.ndarray, n_points: int, seed: int, shape: str = "rect", rings: int = 32,
               sigma: float = 0.02, max_range: float = 100.0) -> np.ndarray:
    """One revolution of a 32-ring spinning LiDAR at map pose (t, q) over the analytic ground with the pit's
    bottom sunk by DEPTH, returned in the sensor's own frame as pcl::PointXYZ records ((n, 4) float32, 1.0f in
    the pad) in firing order (azimuth-major, ring-minor).  Range noise N(0, sigma).  Rays that hit nothing
    within max_range give no point (as a velodyne driver does), so azimuth steps are added until n_points
    returns exist and the list is cut to exactly n_points."""
    rng = np.random.Generator(np.random.PCG64(seed))      # an explicit bit generator: the stream is fixed by NumPy's compatibility policy
    R = _rot(q)
    elev = np.deg2rad(np.linspace(-25.0, 15.0, rings))
    boxes = pit_boxes(shape)
    cx, cy = PIT_CENTER

    def surface(x, y):
        h = ground_height(x, y)
        return np.where(_inside_any(x - cx, y - cy, boxes), h - DEPTH, h)

    out = np.zeros((0, 4), np.float32)
    n_az = max(16, int(np.ceil(n_points / rings * 1.15)))
    for _ in range(6):
        az = (np.arange(n_az) + 0.5) * (2 * np.pi / n_az)
        A2, E2 = np.meshgrid(az, elev, indexing="ij")
        d_s = np.stack([np.cos(E2) * np.cos(A2), np.cos(E2) * np.sin(A2), np.sin(E2)], -1).reshape(-1, 3)
        d_m = d_s @ R.T
        down = d_m[:, 2] < -1e-3
        rng_hit = np.full(d_m.shape[0], np.inf)
        # fixed-point iteration on the height field (gentle slopes: converges in a few rounds)
        tt = np.where(down, (t[2] - 0.0) / np.maximum(-d_m[:, 2], 1e-3), np.inf)
        for _it in range(6):
            ok = np.isfinite(tt)
            x = t[0] + d_m[:, 0] * np.where(ok, tt, 0.0)
            y = t[1] + d_m[:, 1] * np.where(ok, tt, 0.0)
            tt = np.where(ok, (t[2] - surface(x, y)) / np.maximum(-d_m[:, 2], 1e-3), np.inf)
        rng_hit = np.where(down & (tt > 0.3) & (tt < max_range), tt, np.inf)
        hit = np.isfinite(rng_hit)
        r = rng_hit[hit] + rng.normal(0.0, sigma, int(hit.sum()))
        pts = d_s[hit] * r[:, None]
        if pts.shape[0] >= n_points:
            out = np.ones((n_points, 4), np.float32)
            out[:, :3] = pts[:n_points].astype(np.float32)
            return out
        n_az = int(n_az * max(1.2, 1.1 * n_points / max(pts.shape[0], 1)))
    raise RuntimeError("lidar_scan: could not reach the requested number of returns")


@dataclass
class PipelineInputs:
    robot_scan: np.ndarray      # (n, 4) float32 PointXYZ, sensor frame   /four_wheel_robot/velodyne_points
    zx120_scan: np.ndarray      #                                          /zx120/velodyne_points
    robot_tf: tuple             # (t, q) of map -> four_wheel_robot/velodyne_link
    zx120_tf: tuple             # (t, q) of map -> zx120/velodyne_link
    scene: Scene                # the sweep the pipeline feeds (terrain, cells, candidates)


def pipeline_inputs(n_scan: int = 1_000_000, sweep: str = "cfg2", seeds=(4, 5)) -> PipelineInputs:
    """BASELINE config 4: two synthetic n_scan-point 32-ring scans (seeds 4 and 5) that go through
    crop + voxel downsample, are merged into the map frame, and feed a `sweep`-sized visibility sweep whose
    fixed sensor sees the zx120 filter output as its observed cloud (virtual_lidar.cpp:81-85)."""
    sc = build(sweep)
    shape = CONFIGS[sweep]["shape"]
    zt, zq = sensor_in_map(dict(xyz=(0.0, 0.0, 0.0), yaw=0.0), ZX120_VELODYNE)
    rt, rq = sensor_in_map(ROBOT_BASE, ROBOT_VELODYNE)
    return PipelineInputs(robot_scan=lidar_scan(rt, rq, n_scan, seeds[0], shape),
                          zx120_scan=lidar_scan(zt, zq, n_scan, seeds[1], shape),
                          robot_tf=(rt, rq), zx120_tf=(zt, zq), scene=sc)
