"""Thin object wrapper over the plc_* C ABI (include/plc.h).  No computation happens in Python:
every method forwards to libplc.so and raises PlcError on a non-zero status."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _abi as A


@dataclass
class SweepResult:
    best_index: int
    best_score: float
    best_covered: int
    n_cells: int
    fixed_total: float
    counters: dict
    scores: np.ndarray | None = None
    covered: np.ndarray | None = None
    mask: np.ndarray | None = None          # (n_local, mask_words) uint32
    cell_flags: np.ndarray | None = None    # (n_cells,) uint8
    fixed_scores: np.ndarray | None = None  # (n_cells,) float64
    extra: dict = field(default_factory=dict)

    def mask_bits(self) -> np.ndarray:
        """(n_local, n_cells) bool view of the packed visibility mask."""
        bits = np.unpackbits(self.mask.view(np.uint8), axis=1, bitorder="little")
        return bits[:, : self.n_cells].astype(bool)


def _counters(c: A.Counters) -> dict:
    return {k: int(getattr(c, k)) for k, _ in A.Counters._fields_}


class Context:
    """One plc_ctx: one CUDA device, one stream.  Not thread-safe (same contract as the C ABI)."""

    def __init__(self, device: int = 0, stream: int | None = None):
        self._L = A.lib()
        h = C.c_void_p()
        st = self._L.plc_ctx_create(int(device), C.c_void_p(stream) if stream else None, C.byref(h))
        if st != A.PLC_OK:
            raise A.PlcError(st, "plc_ctx_create")
        self._h = h
        self.device = device
        self.n_cells = 0
        self.counting = False        # plc_vl_count_samples: any-hit traversal by default
        self._keep = []

    def close(self):
        for buf in self.__dict__.pop("_graph_pool", {}).values():
            buf.free()
        if getattr(self, "_h", None):
            self._L.plc_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, st: int, where: str):
        if st != A.PLC_OK:
            raise A.PlcError(st, where, self._L.plc_last_error(self._h).decode())

    @property
    def launch_count(self) -> int:
        return int(self._L.plc_launch_count(self._h))

    def sync(self):
        self._check(self._L.plc_ctx_sync(self._h), "plc_ctx_sync")

    # ---- virtual_lidar -------------------------------------------------------------------------
    def set_terrain(self, cloud: np.ndarray, point_step: int | None = None, off=(0, 4, 8)):
        v = A.cloud_view(cloud, point_step, off)
        self._check(self._L.plc_vl_set_terrain(self._h, C.byref(v)), "plc_vl_set_terrain")

    def set_observed(self, cloud: np.ndarray, point_step: int | None = None, off=(0, 4, 8)):
        v = A.cloud_view(cloud, point_step, off)
        self._check(self._L.plc_vl_set_observed(self._h, C.byref(v)), "plc_vl_set_observed")

    def set_cells(self, cells: np.ndarray):
        cells = np.ascontiguousarray(cells, dtype=A.CELL_DTYPE)
        self._check(self._L.plc_vl_set_cells(self._h, A.ptr(cells) if cells.size else None, cells.size),
                    "plc_vl_set_cells")
        self.n_cells = int(cells.size)

    def _alloc_out(self, n_local, want_scores, want_mask, want_flags, want_fixed):
        out = A.SweepOut()
        arrays = {}
        words = (self.n_cells + 31) // 32
        if want_scores:
            arrays["scores"] = np.zeros(n_local, np.float64)
            arrays["covered"] = np.zeros(n_local, np.int32)
        if want_mask:
            arrays["mask"] = np.zeros((n_local, words), np.uint32)
        if want_flags:
            arrays["cell_flags"] = np.zeros(self.n_cells, np.uint8)
        if want_fixed:
            arrays["fixed_scores"] = np.zeros(self.n_cells, np.float64)
        for k, a in arrays.items():
            setattr(out, k, a.ctypes.data)
        return out, arrays

    def _result(self, out: A.SweepOut, arrays: dict) -> SweepResult:
        return SweepResult(int(out.best_index), float(out.best_score), int(out.best_covered),
                           int(out.n_cells), float(out.fixed_total), _counters(out.counters), **arrays)

    def sweep(self, params: A.Params, fixed: np.ndarray, cands: np.ndarray, begin: int = 0,
              end: int | None = None, *, scores=True, mask=False, flags=False, fixed_scores=False) -> SweepResult:
        """plc_vl_sweep: host buffers in, host results out (H2D and D2H inside the call)."""
        cands = np.ascontiguousarray(cands, dtype=A.POSE_DTYPE)
        fixed = np.ascontiguousarray(fixed, dtype=A.POSE_DTYPE).reshape(1)
        end = cands.size if end is None else end
        out, arrays = self._alloc_out(max(end - begin, 0), scores, mask, flags, fixed_scores)
        st = self._L.plc_vl_sweep(self._h, C.byref(params), A.ptr(fixed), A.ptr(cands) if cands.size else None,
                                  cands.size, begin, end, C.byref(out))
        self._check(st, "plc_vl_sweep")
        return self._result(out, arrays)

    def upload_candidates(self, params: A.Params, fixed: np.ndarray, cands: np.ndarray, begin: int = 0,
                          end: int | None = None):
        cands = np.ascontiguousarray(cands, dtype=A.POSE_DTYPE)
        fixed = np.ascontiguousarray(fixed, dtype=A.POSE_DTYPE).reshape(1)
        end = cands.size if end is None else end
        self._n_local = end - begin
        st = self._L.plc_vl_upload_candidates(self._h, C.byref(params), A.ptr(fixed),
                                              A.ptr(cands) if cands.size else None, cands.size, begin, end)
        self._check(st, "plc_vl_upload_candidates")

    def upload_candidates_strided(self, params: A.Params, fixed: np.ndarray, cands: np.ndarray, first: int, stride: int):
        """This context evaluates candidates first, first + stride, ... (round-robin share of the list)."""
        cands = np.ascontiguousarray(cands, dtype=A.POSE_DTYPE)
        fixed = np.ascontiguousarray(fixed, dtype=A.POSE_DTYPE).reshape(1)
        self._n_local = len(range(first, cands.size, stride))
        st = self._L.plc_vl_upload_candidates_strided(self._h, C.byref(params), A.ptr(fixed),
                                                      A.ptr(cands) if cands.size else None, cands.size, first, stride)
        self._check(st, "plc_vl_upload_candidates_strided")

    def sweep_strided(self, params: A.Params, fixed: np.ndarray, cands: np.ndarray, first: int, stride: int, *,
                      scores=True, mask=False, flags=False, fixed_scores=False) -> SweepResult:
        """plc_vl_sweep_strided: arrays are indexed by local candidate i (global index first + i * stride)."""
        cands = np.ascontiguousarray(cands, dtype=A.POSE_DTYPE)
        fixed = np.ascontiguousarray(fixed, dtype=A.POSE_DTYPE).reshape(1)
        out, arrays = self._alloc_out(len(range(first, cands.size, stride)), scores, mask, flags, fixed_scores)
        st = self._L.plc_vl_sweep_strided(self._h, C.byref(params), A.ptr(fixed), A.ptr(cands) if cands.size else None,
                                          cands.size, first, stride, C.byref(out))
        self._check(st, "plc_vl_sweep_strided")
        return self._result(out, arrays)

    def run_async(self, want_mask: bool = False, want_flags: bool = False):
        self._check(self._L.plc_vl_run_async(self._h, int(want_mask), int(want_flags)), "plc_vl_run_async")

    def fetch(self, *, scores=True, mask=False, flags=False, fixed_scores=False) -> SweepResult:
        out, arrays = self._alloc_out(self._n_local, scores, mask, flags, fixed_scores)
        self._check(self._L.plc_vl_fetch(self._h, C.byref(out)), "plc_vl_fetch")
        return self._result(out, arrays)

    def result_device_ptr(self) -> int:
        p = C.c_void_p()
        self._check(self._L.plc_vl_result_device_ptr(self._h, C.byref(p)), "plc_vl_result_device_ptr")
        return int(p.value)

    def last_sweep_kernel_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self._L.plc_vl_last_sweep_kernel_ms(self._h, C.byref(ms)), "plc_vl_last_sweep_kernel_ms")
        return float(ms.value)

    def count_samples(self, on: bool = True):
        """plc_vl_count_samples: counting mode (the reference's evaluated-sample count in `counters["samples"]`)
        instead of the default any-hit traversal."""
        self._check(self._L.plc_vl_count_samples(self._h, 1 if on else 0), "plc_vl_count_samples")
        self.counting = bool(on)

    @property
    def graph_replays(self) -> int:
        """plc_vl_graph_replays: sweeps that were replayed from the context's CUDA graph."""
        n = C.c_uint64(0)
        self._check(self._L.plc_vl_graph_replays(self._h, C.byref(n)), "plc_vl_graph_replays")
        return int(n.value)

    def device_times(self) -> dict:
        """Kernel time (ms) of the last call of each kind, host<->device copies excluded."""
        t = A.DeviceTimes()
        self._check(self._L.plc_last_device_times(self._h, C.byref(t)), "plc_last_device_times")
        return {k: float(getattr(t, k)) for k, _ in A.DeviceTimes._fields_ if k != "reserved"}

    def grid_info(self):
        t, o = A.GridInfo(), A.GridInfo()
        self._check(self._L.plc_vl_grid_info(self._h, C.byref(t), C.byref(o)), "plc_vl_grid_info")
        conv = lambda g: {"n_points": int(g.n_points), "n_voxels": int(g.n_voxels), "grid_bytes": int(g.grid_bytes),
                          "dim": list(g.dim), "origin": list(g.origin), "voxel": float(g.voxel)}
        return conv(t), conv(o)

    # ---- pointcloud_filter / pointcloud_merger -------------------------------------------------
    def filter(self, cloud: np.ndarray, box: A.CropBox, leaf: float, point_step: int | None = None,
               off=(0, 4, 8)):
        """plc_filter: crop box + voxel-grid centroid downsample.  Returns ((n,4) float32, n_cropped)."""
        v = A.cloud_view(cloud, point_step, off)
        cap = max(int(v.n_points), 1)
        out = np.empty((cap, 4), np.float32)      # only [:n_out] is returned; zeroing 16 B per input point costs more than the call
        n_out, n_crop = C.c_uint64(0), C.c_uint64(0)
        st = self._L.plc_filter(self._h, C.byref(v), C.byref(box), float(leaf), out.ctypes.data, cap,
                                C.byref(n_out), C.byref(n_crop))
        self._check(st, "plc_filter")
        return out[: n_out.value], int(n_crop.value)

    def merge(self, sources):
        """plc_merge.  sources: list of dicts {cloud, t(3), q(4 xyzw), rgb(3), present, point_step?, off?}.
        Returns (merged (n,32) uint8 as structured bytes, [per-source arrays])."""
        n = len(sources)
        arr = (A.MergeSrc * n)()
        total = 0
        each = []
        keep = []
        for i, s in enumerate(sources):
            v = A.cloud_view(s["cloud"], s.get("point_step"), s.get("off", (0, 4, 8)))
            keep.append(s["cloud"])
            arr[i].cloud = v
            arr[i].t[:] = list(map(float, s["t"]))
            arr[i].q[:] = list(map(float, s["q"]))
            arr[i].r, arr[i].g, arr[i].b = s["rgb"]
            arr[i].present = 1 if s.get("present", True) else 0
            cnt = int(v.n_points) if arr[i].present else 0
            total += cnt
            each.append(np.zeros((max(cnt, 1), 32), np.uint8))
        merged = np.zeros((max(total, 1), 32), np.uint8)
        ptrs = (C.c_void_p * n)(*[e.ctypes.data for e in each])
        n_each = (C.c_uint64 * n)()
        n_out = C.c_uint64(0)
        st = self._L.plc_merge(self._h, arr, n, merged.ctypes.data, max(total, 1), C.byref(n_out), ptrs, n_each)
        self._check(st, "plc_merge")
        return merged[: n_out.value], [e[: n_each[i]] for i, e in enumerate(each)]


    def excavate(self, matched: np.ndarray, base_t, base_q, params: A.ExcavationParams | None = None,
                 point_step: int | None = None, off=(0, 4, 8), off_rgb: int | None = None):
        """excavated_surface_generator's data path (plc_excavate).  matched: (n, 8) float32 PointXYZRGB records
        (rgb at byte 16) or any layout via point_step/off/off_rgb.  Returns (terrain records (n, 32) uint8,
        area records (m, 32) uint8, info)."""
        if params is None:
            params = A.excavation_params()
        v = A.cloud_view(matched, point_step, off)
        if off_rgb is None:
            off_rgb = 16 if v.point_step >= 20 else 0xFFFFFFFF
        n_surface, n_area = A.excavation_counts(params)
        terrain = np.empty((int(v.n_points) + n_surface, 32), np.uint8)
        area = np.empty((n_area, 32), np.uint8)
        nt, na = C.c_uint64(0), C.c_uint64(0)
        info = A.ExcavationInfo()
        t = np.ascontiguousarray(base_t, np.float64)
        q = np.ascontiguousarray(base_q, np.float64)
        st = self._L.plc_excavate(self._h, C.byref(v), off_rgb, t.ctypes.data, q.ctypes.data, C.byref(params),
                                  terrain.ctypes.data, terrain.shape[0], C.byref(nt), area.ctypes.data, area.shape[0],
                                  C.byref(na), C.byref(info))
        self._check(st, "plc_excavate")
        return terrain[: nt.value], area[: na.value], info

    def drivable_area(self, scan: np.ndarray, t, q, robot_xy, start_xy, params: A.DrivableParams | None = None,
                      point_step: int | None = None, off=(0, 4, 8)):
        """calc_drivable_area's occupancy grid (plc_drivable_area).  Returns ((height, width) int8 grid or None for an
        empty scan, info)."""
        v = A.cloud_view(scan, point_step, off)
        if params is None:
            params = A.DrivableParams()
            self._L.plc_drivable_params_default(C.byref(params))
        tp = None if t is None else np.ascontiguousarray(t, np.float64)
        qp = None if q is None else np.ascontiguousarray(q, np.float64)
        info = A.DrivableInfo()
        w, h = int(params.map_width / params.grid_resolution), int(params.map_height / params.grid_resolution)
        grid = np.zeros(max(w * h, 1), np.int8)
        st = self._L.plc_drivable_area(self._h, C.byref(v), A.ptr(tp), A.ptr(qp), robot_xy[0], robot_xy[1], start_xy[0],
                                       start_xy[1], C.byref(params), grid.ctypes.data, grid.size, C.byref(info))
        self._check(st, "plc_drivable_area")
        if not info.published:
            return None, info
        return grid[: info.n_cells].reshape(info.height, info.width), info

    def drivable_area_static(self, scan_view: A.CloudView, t, q, robot_xy, start_xy, params: A.DrivableParams,
                             out_dev_ptr: int, capacity: int) -> A.DrivableInfo:
        """plc_drivable_area_static: the allocation-free variant.  `scan_view.data` and `out_dev_ptr` are device memory
        (DevBuffer); the grid is complete after sync().  Returns info."""
        tp = None if t is None else np.ascontiguousarray(t, np.float64)
        qp = None if q is None else np.ascontiguousarray(q, np.float64)
        info = A.DrivableInfo()
        st = self._L.plc_drivable_area_static(self._h, C.byref(scan_view), A.ptr(tp), A.ptr(qp), robot_xy[0], robot_xy[1],
                                              start_xy[0], start_xy[1], C.byref(params), out_dev_ptr, capacity,
                                              C.byref(info))
        self._check(st, "plc_drivable_area_static")
        return info

    # ---- node setup on the device: same results as the host_* functions below
    def compute_normals(self, area: np.ndarray, radius: float = 1.5, point_step: int | None = None, off=(0, 4, 8)) -> np.ndarray:
        v = A.cloud_view(area, point_step, off)
        out = np.zeros((int(v.n_points), 3), np.float32)
        self._check(self._L.plc_vl_compute_normals(self._h, C.byref(v), radius, out.ctypes.data), "plc_vl_compute_normals")
        return out

    def generate_cells(self, area: np.ndarray, normals: np.ndarray | None, grid_resolution: float,
                       vertical_layers: int, normal_search_radius: float = 1.5, point_step: int | None = None,
                       off=(0, 4, 8)):
        v = A.cloud_view(area, point_step, off)
        if normals is not None:
            normals = np.ascontiguousarray(normals, np.float32)
        b = A.GridBounds()
        n = C.c_uint64(0)
        # first try with room for the last count (or 64 Ki cells): one pass when it fits, else the count comes back
        # with PLC_E_CAPACITY and the second call is exact
        cells = np.empty(max(getattr(self, "_cells_guess", 0), 1 << 16), A.CELL_DTYPE)
        st = self._L.plc_vl_generate_cells(self._h, C.byref(v), A.ptr(normals), grid_resolution, vertical_layers,
                                           normal_search_radius, cells.ctypes.data, cells.size, C.byref(n), C.byref(b))
        if st == A.PLC_E_CAPACITY:
            cells = np.empty(n.value, A.CELL_DTYPE)
            st = self._L.plc_vl_generate_cells(self._h, C.byref(v), A.ptr(normals), grid_resolution, vertical_layers,
                                               normal_search_radius, cells.ctypes.data, cells.size, C.byref(n), C.byref(b))
        self._check(st, "plc_vl_generate_cells")
        self._cells_guess = int(n.value)
        cells = cells[: n.value]          # a view: copying a structured array goes field by field
        return cells, b

    def generate_candidates(self, terrain: np.ndarray | None, bounds: A.GridBounds, fixed: np.ndarray,
                            num_candidates: int, search_radius: float, sensor_height: float,
                            point_step: int | None = None, off=(0, 4, 8)) -> np.ndarray:
        """terrain=None: the cloud last given to set_terrain (already on the device)."""
        v = A.cloud_view(terrain, point_step, off) if terrain is not None else None
        fixed = np.ascontiguousarray(fixed, dtype=A.POSE_DTYPE).reshape(1)
        side = int(np.ceil(np.sqrt(max(num_candidates, 1))))
        out = np.zeros(side * side, A.POSE_DTYPE)
        n = C.c_uint64(0)
        st = self._L.plc_vl_generate_candidates(self._h, C.byref(v) if v is not None else None, C.byref(bounds),
                                                fixed.ctypes.data, num_candidates, search_radius, sensor_height,
                                                out.ctypes.data, out.size, C.byref(n))
        self._check(st, "plc_vl_generate_candidates")
        return out[: n.value].copy()


class DevBuffer:
    """plc_dev_alloc: device memory a caller hands from one stage to the next."""

    def __init__(self, ctx: "Context", nbytes: int):
        self._ctx, self.nbytes = ctx, int(nbytes)
        p = C.c_void_p()
        ctx._check(ctx._L.plc_dev_alloc(ctx._h, self.nbytes, C.byref(p)), "plc_dev_alloc")
        self.ptr = p.value

    def free(self):
        if getattr(self, "ptr", None) and getattr(self._ctx, "_h", None):
            self._ctx._L.plc_dev_free(self._ctx._h, self.ptr)
        self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def upload(self, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        self._ctx._check(self._ctx._L.plc_dev_upload(self._ctx._h, self.ptr, arr.ctypes.data, arr.nbytes), "plc_dev_upload")
        return self

    def download(self, nbytes: int, dtype=np.uint8) -> np.ndarray:
        out = np.empty(int(nbytes) // np.dtype(dtype).itemsize, dtype)
        self._ctx._check(self._ctx._L.plc_dev_download(self._ctx._h, out.ctypes.data, self.ptr, out.nbytes), "plc_dev_download")
        return out

    def view(self, n_points: int, point_step: int, off=(0, 4, 8)) -> A.CloudView:
        v = A.CloudView()
        v.data = self.ptr if n_points else None
        v.n_points, v.point_step = int(n_points), int(point_step)
        v.off_x, v.off_y, v.off_z = off
        return v

    def export(self) -> bytes:
        """plc_dev_export: 64 opaque bytes another process on the same GPU turns back into this buffer."""
        h = (C.c_ubyte * 64)()
        self._ctx._check(self._ctx._L.plc_dev_export(self._ctx._h, self.ptr, h), "plc_dev_export")
        return bytes(h)


class ImportedBuffer(DevBuffer):
    """plc_dev_import: a buffer another process exported; released (unmapped), never freed, by this side."""

    def __init__(self, ctx: "Context", handle: bytes, nbytes: int):
        assert len(handle) == 64
        self._ctx, self.nbytes = ctx, int(nbytes)
        p = C.c_void_p()
        h = (C.c_ubyte * 64).from_buffer_copy(handle)
        ctx._check(ctx._L.plc_dev_import(ctx._h, h, C.byref(p)), "plc_dev_import")
        self.ptr = p.value

    def free(self):
        if getattr(self, "ptr", None) and getattr(self._ctx, "_h", None):
            self._ctx._L.plc_dev_release(self._ctx._h, self.ptr)
        self.ptr = None

    def export(self) -> bytes:
        raise A.PlcError(A.PLC_E_INVALID_ARG, "an imported buffer is exported by its owner")


# ---- host-side setup (include/plc_host.h) -------------------------------------------------------
def host_generate_cells(area: np.ndarray, normals: np.ndarray | None, grid_resolution: float,
                        vertical_layers: int, normal_search_radius: float = 1.5):
    L = A.lib()
    v = A.cloud_view(area)
    if normals is not None:
        normals = np.ascontiguousarray(normals, np.float32)
    b = A.GridBounds()
    n = C.c_uint64(0)
    st = L.plc_host_generate_cells(C.byref(v), A.ptr(normals), grid_resolution, vertical_layers,
                                   normal_search_radius, None, 0, C.byref(n), C.byref(b))
    if st not in (A.PLC_OK, A.PLC_E_CAPACITY):
        raise A.PlcError(st, "plc_host_generate_cells")
    cells = np.zeros(n.value, A.CELL_DTYPE)
    if n.value:
        st = L.plc_host_generate_cells(C.byref(v), A.ptr(normals), grid_resolution, vertical_layers,
                                       normal_search_radius, cells.ctypes.data, cells.size, C.byref(n), C.byref(b))
        if st != A.PLC_OK:
            raise A.PlcError(st, "plc_host_generate_cells")
    return cells, b


def host_compute_normals(area: np.ndarray, radius: float = 1.5) -> np.ndarray:
    L = A.lib()
    v = A.cloud_view(area)
    out = np.zeros((area.shape[0], 3), np.float32)
    st = L.plc_host_compute_normals(C.byref(v), radius, out.ctypes.data)
    if st != A.PLC_OK:
        raise A.PlcError(st, "plc_host_compute_normals")
    return out


def host_generate_candidates(terrain: np.ndarray, bounds: A.GridBounds, fixed: np.ndarray,
                             num_candidates: int, search_radius: float, sensor_height: float) -> np.ndarray:
    L = A.lib()
    v = A.cloud_view(terrain)
    fixed = np.ascontiguousarray(fixed, dtype=A.POSE_DTYPE).reshape(1)
    side = int(np.ceil(np.sqrt(num_candidates)))
    out = np.zeros(side * side, A.POSE_DTYPE)
    n = C.c_uint64(0)
    st = L.plc_host_generate_candidates(C.byref(v), C.byref(bounds), fixed.ctypes.data, num_candidates,
                                        search_radius, sensor_height, out.ctypes.data, out.size, C.byref(n))
    if st != A.PLC_OK:
        raise A.PlcError(st, "plc_host_generate_candidates")
    return out[: n.value].copy()


class Group:
    """plc_group_*: the sweep on several GPUs from one process (terrain and cells replicated, candidate poses
    sharded, one NCCL all-gather of 16 bytes per device).  Results are indexed by global candidate."""

    def __init__(self, devices):
        self._L = A.lib()
        devs = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        st = self._L.plc_group_create(devs, len(devices), C.byref(h))
        if st != A.PLC_OK:
            raise A.PlcError(st, "plc_group_create", "")
        self._h = h
        self.devices = list(devices)
        self.n_cells = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.plc_group_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st, where):
        if st != A.PLC_OK:
            raise A.PlcError(st, where, self._L.plc_group_last_error(self._h).decode())

    def set_terrain(self, cloud, point_step=None, off=(0, 4, 8)):
        v = A.cloud_view(cloud, point_step, off)
        self._check(self._L.plc_group_set_terrain(self._h, C.byref(v)), "plc_group_set_terrain")

    def set_observed(self, cloud, point_step=None, off=(0, 4, 8)):
        v = A.cloud_view(cloud, point_step, off)
        self._check(self._L.plc_group_set_observed(self._h, C.byref(v)), "plc_group_set_observed")

    def set_cells(self, cells):
        cells = np.ascontiguousarray(cells, dtype=A.CELL_DTYPE)
        self._check(self._L.plc_group_set_cells(self._h, A.ptr(cells) if cells.size else None, cells.size),
                    "plc_group_set_cells")
        self.n_cells = int(cells.size)

    def sweep(self, params, fixed, cands, *, scores=True, mask=False, flags=False, fixed_scores=False) -> SweepResult:
        cands = np.ascontiguousarray(cands, dtype=A.POSE_DTYPE)
        fixed = np.ascontiguousarray(fixed, dtype=A.POSE_DTYPE).reshape(1)
        out, arrays = Context._alloc_out(self, cands.size, scores, mask, flags, fixed_scores)
        st = self._L.plc_group_sweep(self._h, C.byref(params), A.ptr(fixed), A.ptr(cands) if cands.size else None,
                                     cands.size, C.byref(out))
        self._check(st, "plc_group_sweep")
        return Context._result(self, out, arrays)
