"""Headless drivers of the ROS-free node cores (include/plc_node.h) — what the rclcpp wrappers in
ros2/ call, minus the message plumbing.  Used by tests and tools; no computation happens here."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi as A
from .context import Context


class VirtualLidarNode:
    """plc_vlnode: state and callbacks of SimplifiedDualLidarOptimizer (virtual_lidar.cpp:60-548)."""

    def __init__(self, ctx):
        """ctx: a Context (one GPU) or a context.Group (several GPUs, sweeps sharded over them)."""
        self._L = A.lib()
        self._ctx = ctx
        self._on_group = not isinstance(ctx, Context)
        h = C.c_void_p()
        create = self._L.plc_vlnode_create_on_group if self._on_group else self._L.plc_vlnode_create
        st = create(ctx._h, C.byref(h))
        if st != A.PLC_OK:
            raise A.PlcError(st, "plc_vlnode_create")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.plc_vlnode_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st, where):
        if st != A.PLC_OK:
            err = self._L.plc_group_last_error if self._on_group else self._L.plc_last_error
            raise A.PlcError(st, where, err(self._ctx._h).decode())

    def set_parameter(self, name: str, value):
        if isinstance(value, (int, np.integer)) and not isinstance(value, bool):
            st = self._L.plc_vlnode_set_parameter_int(self._h, name.encode(), int(value))
        else:
            st = self._L.plc_vlnode_set_parameter_double(self._h, name.encode(), float(value))
        self._check(st, f"set_parameter({name})")

    def update_parameters(self):
        """updateParameters() as the reference's constructor calls it (:73): launch-time overrides hold from the first message."""
        self._check(self._L.plc_vlnode_update_parameters(self._h), "update_parameters")

    def excavation_area_callback(self, cloud, point_step=None, off=(0, 4, 8)):
        v = A.cloud_view(cloud, point_step, off)
        self._check(self._L.plc_vlnode_excavation_area_callback(self._h, C.byref(v)), "excavation_area_callback")

    def terrain_callback(self, cloud, point_step=None, off=(0, 4, 8)):
        v = A.cloud_view(cloud, point_step, off)
        self._check(self._L.plc_vlnode_terrain_callback(self._h, C.byref(v)), "terrain_callback")

    def zx120_points_callback(self, cloud, point_step=None, off=(0, 4, 8)):
        v = A.cloud_view(cloud, point_step, off)
        self._check(self._L.plc_vlnode_zx120_points_callback(self._h, C.byref(v)), "zx120_points_callback")

    def set_base_transform(self, xyz=None):
        if xyz is None:
            st = self._L.plc_vlnode_set_base_transform(self._h, 0, 0.0, 0.0, 0.0)
        else:
            st = self._L.plc_vlnode_set_base_transform(self._h, 1, float(xyz[0]), float(xyz[1]), float(xyz[2]))
        self._check(st, "set_base_transform")

    def run_optimization(self) -> A.NodeResult:
        out = A.NodeResult()
        self._check(self._L.plc_vlnode_run_optimization(self._h, C.byref(out)), "run_optimization")
        return out

    def candidates(self) -> np.ndarray:
        n = C.c_uint64(0)
        self._L.plc_vlnode_get_candidates(self._h, None, 0, C.byref(n))
        out = np.zeros(n.value, A.POSE_DTYPE)
        if n.value:
            self._check(self._L.plc_vlnode_get_candidates(self._h, out.ctypes.data, out.size, C.byref(n)), "get_candidates")
        return out

    def cells(self):
        n = C.c_uint64(0)
        self._L.plc_vlnode_get_cells(self._h, None, None, 0, C.byref(n))
        cells, flags = np.zeros(n.value, A.CELL_DTYPE), np.zeros(n.value, np.uint8)
        if n.value:
            self._check(self._L.plc_vlnode_get_cells(self._h, cells.ctypes.data, flags.ctypes.data, cells.size,
                                                     C.byref(n)), "get_cells")
        return cells, flags

    def bounds(self) -> A.GridBounds:
        b = A.GridBounds()
        self._check(self._L.plc_vlnode_get_bounds(self._h, C.byref(b)), "get_bounds")
        return b

    def take_log(self) -> list[str]:
        text = self._L.plc_vlnode_take_log(self._h).decode()
        return text.split("\n")[:-1] if text else []


def filter_scan(ctx: Context, scan, front=15.0, side=10.0, height=10.0, leaf=0.2, point_step=None, off=(0, 4, 8)):
    """plc_filter_scan: pointcloud_filter's processCloudSimple.  Returns (n, 4) float32 (PointXYZ wire records)."""
    L = A.lib()
    v = A.cloud_view(scan, point_step, off)
    cap = max(int(v.n_points), 1)
    out = np.empty((cap, 4), np.float32)
    n = C.c_uint64(0)
    st = L.plc_filter_scan(ctx._h, C.byref(v), front, side, height, leaf, out.ctypes.data, cap, C.byref(n))
    if st != A.PLC_OK:
        raise A.PlcError(st, "plc_filter_scan", L.plc_last_error(ctx._h).decode())
    return out[: n.value]


def merge_clouds(ctx: Context, robot: dict | None, backhoe: dict | None):
    """plc_merge_clouds: pointcloud_merger's processPointClouds.  Each source is None (no message yet) or
    {cloud (n,k) float32, t (3) or None for a TF failure, q (4 xyzw)}.  Returns (merged, robot, backhoe) as
    (n, 32) uint8 PointXYZRGB wire records."""
    L = A.lib()
    ins, keep, caps = [], [], []
    for src in (robot, backhoe):
        m = A.MergerInput()
        if src is not None:
            v = A.cloud_view(src["cloud"], src.get("point_step"), src.get("off", (0, 4, 8)))
            keep.append(src["cloud"])
            m.cloud = v
            m.has_cloud = 1
            if src.get("t") is not None:
                m.has_transform = 1
                m.t[:] = list(map(float, src["t"]))
                m.q[:] = list(map(float, src["q"]))
            caps.append(int(v.n_points))
        else:
            caps.append(0)
        ins.append(m)
    bufs = [np.zeros((max(sum(caps), 1), 32), np.uint8), np.zeros((max(caps[0], 1), 32), np.uint8),
            np.zeros((max(caps[1], 1), 32), np.uint8)]
    ns = [C.c_uint64(0) for _ in range(3)]
    st = L.plc_merge_clouds(ctx._h, C.byref(ins[0]), C.byref(ins[1]),
                            bufs[0].ctypes.data, bufs[0].shape[0], C.byref(ns[0]),
                            bufs[1].ctypes.data, bufs[1].shape[0], C.byref(ns[1]),
                            bufs[2].ctypes.data, bufs[2].shape[0], C.byref(ns[2]))
    if st != A.PLC_OK:
        raise A.PlcError(st, "plc_merge_clouds", L.plc_last_error(ctx._h).decode())
    return tuple(b[: n.value] for b, n in zip(bufs, ns))


def run_pipeline(ctx: Context, inp, leaf: float = 0.2, **sweep_kw):
    """BASELINE config 4: the launch file's dataflow for one scan pair:
    pointcloud_filter (both scans) -> pointcloud_merger -> virtual_lidar, where the zx120 filter output is
    also the fixed sensor's observed cloud (virtual_lidar.cpp:81-85; sensor frame, as upstream).
    Returns a dict with every intermediate product and the kernel time of each stage."""
    sc = inp.scene
    out = {}
    times = {}
    out["robot_filtered"] = filter_scan(ctx, inp.robot_scan, leaf=leaf)
    times["filter_robot_ms"] = ctx.device_times()["filter_ms"]
    out["zx120_filtered"] = filter_scan(ctx, inp.zx120_scan, leaf=leaf)
    times["filter_zx120_ms"] = ctx.device_times()["filter_ms"]
    merged, rc, bc = merge_clouds(ctx, dict(cloud=out["robot_filtered"], t=inp.robot_tf[0], q=inp.robot_tf[1]),
                                  dict(cloud=out["zx120_filtered"], t=inp.zx120_tf[0], q=inp.zx120_tf[1]))
    times["merge_ms"] = ctx.device_times()["merge_ms"]
    out.update(merged=merged, robot_coloured=rc, backhoe_coloured=bc)
    ctx.set_terrain(sc.terrain)
    times["set_terrain_ms"] = ctx.device_times()["set_terrain_ms"]
    ctx.set_cells(sc.cells)
    ctx.set_observed(out["zx120_filtered"], point_step=16)
    times["set_observed_ms"] = ctx.device_times()["set_observed_ms"]
    out["sweep"] = ctx.sweep(sc.params, sc.fixed, sc.candidates, **sweep_kw)
    times["sweep_kernel_ms"] = ctx.device_times()["sweep_kernel_ms"]
    out["device_ms"] = times
    return out


def run_launch_graph(ctx: Context, inp, leaf: float = 0.2, num_candidates: int = 100, grid_resolution: float = 0.1,
                     vertical_layers: int = 10, sensor_height: float = 1.1, search_radius: float = 3.0,
                     excavation: A.ExcavationParams | None = None, robot_base_xy=None, **sweep_kw):
    """The whole dataflow of launch/start.launch.xml for one scan pair, every data-parallel stage on the device:
    pointcloud_filter (both scans) -> pointcloud_merger -> excavated_surface_generator -> virtual_lidar
    (normals, evaluation cells, candidate poses, sweep), with calc_drivable_area beside it on the robot's
    filtered scan.  Unlike run_pipeline, the sweep's terrain and evaluation area are the generator's outputs for
    the merged cloud, not a synthetic scene.  zx120/base_link sits at the map origin (pointcloud_merger.cpp:218-228).
    Returns every intermediate product and the kernel time of each stage."""
    from .scenes import fixed_pose, ROBOT_BASE
    out, times = {}, {}
    out["robot_filtered"] = filter_scan(ctx, inp.robot_scan, leaf=leaf)
    times["filter_robot_ms"] = ctx.device_times()["filter_ms"]
    out["zx120_filtered"] = filter_scan(ctx, inp.zx120_scan, leaf=leaf)
    times["filter_zx120_ms"] = ctx.device_times()["filter_ms"]
    merged, rc, bc = merge_clouds(ctx, dict(cloud=out["robot_filtered"], t=inp.robot_tf[0], q=inp.robot_tf[1]),
                                  dict(cloud=out["zx120_filtered"], t=inp.zx120_tf[0], q=inp.zx120_tf[1]))
    times["merge_ms"] = ctx.device_times()["merge_ms"]
    out.update(merged=merged, robot_coloured=rc, backhoe_coloured=bc)
    # /matched_point_cloud -> /excavated_terrain + /excavation_area (PointXYZRGB records, rgb at byte 16)
    base_t, base_q = np.zeros(3), np.array([0.0, 0.0, 0.0, 1.0])
    terrain, area, info = ctx.excavate(merged.reshape(-1), base_t, base_q, excavation, point_step=32, off_rgb=16)
    times["excavate_ms"] = ctx.device_times()["excavate_ms"]
    out.update(terrain=terrain, area=area, excavation_info=info)
    # virtual_lidar: excavationAreaCallback, terrainCallback, zx120PointsCallback, runOptimization
    normals = ctx.compute_normals(area.reshape(-1), 1.5, point_step=32)
    times["normals_ms"] = ctx.device_times()["normals_ms"]
    cells, bounds = ctx.generate_cells(area.reshape(-1), normals, grid_resolution, vertical_layers, 1.5, point_step=32)
    times["cells_ms"] = ctx.device_times()["cells_ms"]
    ctx.set_terrain(terrain.reshape(-1), point_step=32)
    times["set_terrain_ms"] = ctx.device_times()["set_terrain_ms"]
    ctx.set_cells(cells)
    ctx.set_observed(out["zx120_filtered"], point_step=16)
    times["set_observed_ms"] = ctx.device_times()["set_observed_ms"]
    fixed = fixed_pose()
    cands = ctx.generate_candidates(None, bounds, fixed, num_candidates, search_radius, sensor_height)
    times["candidates_ms"] = ctx.device_times()["candidates_ms"]
    out.update(normals=normals, cells=cells, bounds=bounds, candidates=cands, fixed=fixed)
    out["sweep"] = ctx.sweep(A.default_params(), fixed, cands, **sweep_kw)
    times["sweep_kernel_ms"] = ctx.device_times()["sweep_kernel_ms"]
    # calc_drivable_area on the robot's filtered scan (its TF into the map is the merger's)
    rxy = robot_base_xy if robot_base_xy is not None else ROBOT_BASE["xyz"][:2]
    grid, ginfo = ctx.drivable_area(out["robot_filtered"], inp.robot_tf[0], inp.robot_tf[1], rxy, rxy, point_step=16)
    times["drivable_ms"] = ctx.device_times()["drivable_ms"]
    out.update(occupancy_grid=grid, occupancy_info=ginfo)
    out["device_ms"] = times
    return out


def run_launch_graph_resident(ctx: Context, inp, leaf: float = 0.2, num_candidates: int = 100, grid_resolution: float = 0.1,
                              vertical_layers: int = 10, sensor_height: float = 1.1, search_radius: float = 3.0,
                              excavation: A.ExcavationParams | None = None, robot_base_xy=None, download: bool = False,
                              **sweep_kw):
    """run_launch_graph with every cloud between the stages left in HBM: the two scans come from the
    host (they are sensor input), the evaluation cells, the poses and the results go to the host (the node needs
    them there); filtered scans, merged cloud, excavated terrain, excavation area and normals only ever exist in
    device buffers (plc_dev_alloc) that the next stage reads in place.  `download=True` also copies those
    intermediate clouds back (after the timed part) so that tests can compare them."""
    import time
    from .context import DevBuffer as _DevBuffer
    from .scenes import fixed_pose, ROBOT_BASE
    L = A.lib()
    prm = excavation if excavation is not None else A.excavation_params()
    chk = ctx._check
    out, n = {}, {}
    # a node keeps its device buffers from one message to the next: grow-only pool on the context
    pool = ctx.__dict__.setdefault("_graph_pool", {})

    def DevBuffer(_ctx, nbytes, _name=[0]):
        _name[0] += 1
        key = _name[0]
        buf = pool.get(key)
        if buf is None or buf.nbytes < nbytes:
            if buf is not None:
                buf.free()
            buf = pool[key] = _DevBuffer(ctx, int(nbytes * 1.25) + 256)
        return buf
    t0 = time.perf_counter()
    # pointcloud_filter x2
    filt = {}
    for name, scan in (("robot", inp.robot_scan), ("zx120", inp.zx120_scan)):
        v = A.cloud_view(scan)
        filt[name] = DevBuffer(ctx, max(int(v.n_points), 1) * 16)
        cnt = C.c_uint64(0)
        chk(L.plc_filter_scan(ctx._h, C.byref(v), 15.0, 10.0, 10.0, leaf, filt[name].ptr, int(v.n_points), C.byref(cnt)), "plc_filter_scan")
        n[name] = cnt.value
    # pointcloud_merger
    ins = []
    for name, tf in (("robot", inp.robot_tf), ("zx120", inp.zx120_tf)):
        m = A.MergerInput()
        m.cloud = filt[name].view(n[name], 16)
        m.has_cloud = m.has_transform = 1
        m.t[:] = list(map(float, tf[0]))
        m.q[:] = list(map(float, tf[1]))
        ins.append(m)
    n_m = n["robot"] + n["zx120"]
    merged, own_r, own_b = DevBuffer(ctx, max(n_m, 1) * 32), DevBuffer(ctx, max(n["robot"], 1) * 32), DevBuffer(ctx, max(n["zx120"], 1) * 32)
    c0, c1, c2 = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
    chk(L.plc_merge_clouds(ctx._h, C.byref(ins[0]), C.byref(ins[1]), merged.ptr, n_m, C.byref(c0), own_r.ptr, n["robot"], C.byref(c1),
                           own_b.ptr, n["zx120"], C.byref(c2)), "plc_merge_clouds")
    n["merged"] = c0.value
    # excavated_surface_generator
    ns, na = A.excavation_counts(prm)
    terrain, area = DevBuffer(ctx, (n["merged"] + ns) * 32), DevBuffer(ctx, max(na, 1) * 32)
    nt, nar = C.c_uint64(0), C.c_uint64(0)
    info = A.ExcavationInfo()
    base_t, base_q = np.zeros(3), np.array([0.0, 0.0, 0.0, 1.0])
    vm = merged.view(n["merged"], 32)
    chk(L.plc_excavate(ctx._h, C.byref(vm), 16, base_t.ctypes.data, base_q.ctypes.data, C.byref(prm), terrain.ptr,
                       n["merged"] + ns, C.byref(nt), area.ptr, na, C.byref(nar), C.byref(info)), "plc_excavate")
    n["terrain"], n["area"] = nt.value, nar.value
    # virtual_lidar: normals and cells from the area cloud, indices from the terrain and the zx120 scan, poses, sweep
    va, vt = area.view(n["area"], 32), terrain.view(n["terrain"], 32)
    normals = DevBuffer(ctx, max(n["area"], 1) * 12)
    chk(L.plc_vl_compute_normals(ctx._h, C.byref(va), 1.5, normals.ptr), "plc_vl_compute_normals")
    b = A.GridBounds()
    cnt = C.c_uint64(0)
    cells = np.empty(max(getattr(ctx, "_cells_guess", 0), 1 << 16), A.CELL_DTYPE)   # one pass when the last count fits
    st = L.plc_vl_generate_cells(ctx._h, C.byref(va), normals.ptr, grid_resolution, vertical_layers, 1.5, cells.ctypes.data,
                                 cells.size, C.byref(cnt), C.byref(b))
    if st == A.PLC_E_CAPACITY:
        cells = np.empty(cnt.value, A.CELL_DTYPE)
        st = L.plc_vl_generate_cells(ctx._h, C.byref(va), normals.ptr, grid_resolution, vertical_layers, 1.5, cells.ctypes.data,
                                     cells.size, C.byref(cnt), C.byref(b))
    chk(st, "plc_vl_generate_cells")
    ctx._cells_guess = int(cnt.value)
    cells = cells[: cnt.value]
    chk(L.plc_vl_set_terrain(ctx._h, C.byref(vt)), "plc_vl_set_terrain")
    ctx.set_cells(cells)
    vo = filt["zx120"].view(n["zx120"], 16)
    chk(L.plc_vl_set_observed(ctx._h, C.byref(vo)), "plc_vl_set_observed")
    fixed = fixed_pose()
    cands = ctx.generate_candidates(None, b, fixed, num_candidates, search_radius, sensor_height)
    out["sweep"] = ctx.sweep(A.default_params(), fixed, cands, **sweep_kw)
    # calc_drivable_area
    rxy = robot_base_xy if robot_base_xy is not None else ROBOT_BASE["xyz"][:2]
    dprm = A.DrivableParams()
    L.plc_drivable_params_default(C.byref(dprm))
    dinfo = A.DrivableInfo()
    w, h = int(dprm.map_width / dprm.grid_resolution), int(dprm.map_height / dprm.grid_resolution)
    grid = np.zeros(max(w * h, 1), np.int8)
    vr = filt["robot"].view(n["robot"], 16)
    t_r, q_r = np.ascontiguousarray(inp.robot_tf[0], np.float64), np.ascontiguousarray(inp.robot_tf[1], np.float64)
    chk(L.plc_drivable_area(ctx._h, C.byref(vr), t_r.ctypes.data, q_r.ctypes.data, rxy[0], rxy[1], rxy[0], rxy[1], C.byref(dprm),
                            grid.ctypes.data, grid.size, C.byref(dinfo)), "plc_drivable_area")
    out["wall_ms"] = (time.perf_counter() - t0) * 1e3
    out.update(cells=cells, bounds=b, candidates=cands, fixed=fixed, excavation_info=info, counts=dict(n),
               occupancy_grid=grid[: dinfo.n_cells].reshape(dinfo.height, dinfo.width) if dinfo.published else None)
    if download:
        out["robot_filtered"] = filt["robot"].download(n["robot"] * 16, np.float32).reshape(-1, 4)
        out["zx120_filtered"] = filt["zx120"].download(n["zx120"] * 16, np.float32).reshape(-1, 4)
        out["merged"] = merged.download(n["merged"] * 32).reshape(-1, 32)
        out["terrain"] = terrain.download(n["terrain"] * 32).reshape(-1, 32)
        out["area"] = area.download(n["area"] * 32).reshape(-1, 32)
        out["normals"] = normals.download(n["area"] * 12, np.float32).reshape(-1, 3)
    return out
