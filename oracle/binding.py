"""ctypes binding of the CPU oracle (oracle/oracle.h).  TEST INFRASTRUCTURE: imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs — never by the
product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

from pointcloud_processor_b200 import _abi as A   # shared POD layouts only (include/plc.h)

_HERE = Path(__file__).resolve().parent
# PLC_ORACLE_LIB: another build of the same sources (the ASan/UBSan one of `make -C oracle asan`)
LIB_PATH = Path(os.environ.get("PLC_ORACLE_LIB") or _HERE / "_build" / "liboracle.so")

BRUTE, GRID, KDTREE = 0, 1, 2


class SweepArgs(C.Structure):
    _fields_ = [("terrain", C.c_void_p), ("observed", C.c_void_p), ("cells", C.c_void_p),
                ("n_cells", C.c_uint64), ("prm", C.POINTER(A.Params)), ("fixed", C.c_void_p),
                ("cands", C.c_void_p), ("n_cands", C.c_uint64), ("faithful", C.c_int32),
                ("threads", C.c_int32), ("scores", C.c_void_p), ("covered", C.c_void_p),
                ("mask", C.c_void_p), ("cell_flags", C.c_void_p), ("fixed_scores", C.c_void_p),
                ("best_index", C.c_int64), ("best_score", C.c_double), ("best_covered", C.c_int32),
                ("fixed_total", C.c_double), ("counters", A.Counters)]


_lib = None


def build(force: bool = False):
    if force or not LIB_PATH.exists():
        subprocess.check_call(["make", "-C", os.fspath(_HERE)] + (["-B"] if force else []),
                              stdout=subprocess.DEVNULL)


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(os.fspath(LIB_PATH))
    vp, u64, i64, dbl, f32 = C.c_void_p, C.c_uint64, C.c_int64, C.c_double, C.c_float
    L.orc_index_create.restype = vp
    L.orc_index_create.argtypes = [vp, u64, C.c_int]
    L.orc_index_destroy.argtypes = [vp]
    L.orc_index_destroy.restype = None
    L.orc_index_size.restype = u64
    L.orc_index_size.argtypes = [vp]
    L.orc_index_radius_count.restype = i64
    L.orc_index_radius_count.argtypes = [vp, f32, f32, f32, dbl]
    L.orc_unpack_cloud.argtypes = [C.POINTER(A.CloudView), vp]
    L.orc_unpack_cloud.restype = None
    L.orc_vl_sweep.argtypes = [C.POINTER(SweepArgs)]
    L.orc_generate_cells.restype = i64
    L.orc_generate_cells.argtypes = [vp, u64, vp, dbl, C.c_int32, dbl, vp, u64, C.POINTER(A.GridBounds)]
    L.orc_compute_normals.argtypes = [vp, u64, dbl, vp]
    L.orc_compute_normals.restype = None
    L.orc_ground_height.restype = dbl
    L.orc_ground_height.argtypes = [vp, vp, dbl, dbl]
    L.orc_generate_candidates.restype = i64
    L.orc_generate_candidates.argtypes = [vp, vp, C.POINTER(A.GridBounds), vp, C.c_int32, dbl, dbl, vp, u64]
    L.orc_crop.restype = i64
    L.orc_crop.argtypes = [vp, u64, C.POINTER(A.CropBox), vp, vp]
    L.orc_voxel_grid.restype = i64
    L.orc_voxel_grid.argtypes = [vp, u64, dbl, vp, vp, vp, vp]
    L.orc_transform.argtypes = [vp, u64, vp, vp, vp]
    L.orc_transform.restype = None
    L.orc_excavate.restype = None
    L.orc_excavate.argtypes = [vp, vp, u64, vp, vp, C.POINTER(A.ExcavationParams), vp, C.POINTER(u64), vp, C.POINTER(u64),
                               C.POINTER(A.ExcavationInfo)]
    L.orc_excavation_frame.restype = None
    L.orc_excavation_frame.argtypes = [vp, vp, dbl, dbl, vp, vp]
    L.orc_drivable_area.restype = i64
    L.orc_drivable_area.argtypes = [vp, u64, vp, vp, dbl, dbl, dbl, dbl, dbl, dbl, dbl, dbl, C.c_int, dbl, vp,
                                    C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    _lib = L
    return L


class Index:
    """Radius index over packed float32 xyz (non-finite points dropped)."""

    def __init__(self, xyz: np.ndarray, kind: int = GRID):
        self.xyz = np.ascontiguousarray(xyz[:, :3], np.float32)
        self._L = lib()
        self.h = self._L.orc_index_create(self.xyz.ctypes.data if self.xyz.size else None, self.xyz.shape[0], kind)

    def __del__(self):
        if getattr(self, "h", None):
            self._L.orc_index_destroy(self.h)
            self.h = None

    def size(self) -> int:
        return int(self._L.orc_index_size(self.h))

    def count(self, q, radius: float) -> int:
        return int(self._L.orc_index_radius_count(self.h, float(q[0]), float(q[1]), float(q[2]), float(radius)))


def sweep(terrain: Index | None, observed: Index | None, cells, params, fixed, cands, *, faithful=False,
          threads=1, want_mask=True, flags_inout: np.ndarray | None = None):
    """Returns a dict with scores, covered, mask (C, words), flags, fixed_scores, best_*, counters."""
    L = lib()
    cells = np.ascontiguousarray(cells, A.CELL_DTYPE)
    cands = np.ascontiguousarray(cands, A.POSE_DTYPE)
    fixed = np.ascontiguousarray(fixed, A.POSE_DTYPE).reshape(1)
    M, Cn = cells.size, cands.size
    W = (M + 31) // 32
    a = SweepArgs()
    a.terrain = terrain.h if terrain is not None else None
    a.observed = observed.h if observed is not None else None
    a.cells = cells.ctypes.data if M else None
    a.n_cells = M
    a.prm = C.pointer(params)
    a.fixed = fixed.ctypes.data
    a.cands = cands.ctypes.data if Cn else None
    a.n_cands = Cn
    a.faithful = int(faithful)
    a.threads = int(threads)
    scores = np.zeros(Cn, np.float64)
    covered = np.zeros(Cn, np.int32)
    mask = np.zeros((Cn, W), np.uint32) if want_mask else None
    flags = (flags_inout if flags_inout is not None else np.zeros(M, np.uint8)) if threads <= 1 else None
    fixed_scores = np.zeros(M, np.float64)
    a.scores = scores.ctypes.data
    a.covered = covered.ctypes.data
    a.mask = mask.ctypes.data if mask is not None else None
    a.cell_flags = flags.ctypes.data if flags is not None else None
    a.fixed_scores = fixed_scores.ctypes.data
    rc = L.orc_vl_sweep(C.byref(a))
    if rc != 0:
        raise RuntimeError(f"orc_vl_sweep failed: {rc}")
    return dict(scores=scores, covered=covered, mask=mask, flags=flags, fixed_scores=fixed_scores,
                best_index=int(a.best_index), best_score=float(a.best_score), best_covered=int(a.best_covered),
                fixed_total=float(a.fixed_total),
                counters={k: int(getattr(a.counters, k)) for k, _ in A.Counters._fields_})


def generate_cells(area, normals, res, layers, normal_radius=1.5):
    L = lib()
    area = np.ascontiguousarray(area[:, :3], np.float32)
    nrm = None if normals is None else np.ascontiguousarray(normals, np.float32)
    b = A.GridBounds()
    n = L.orc_generate_cells(area.ctypes.data, area.shape[0], A.ptr(nrm), res, layers, normal_radius, None, 0, C.byref(b))
    out = np.zeros(max(n, 0), A.CELL_DTYPE)
    if n > 0:
        L.orc_generate_cells(area.ctypes.data, area.shape[0], A.ptr(nrm), res, layers, normal_radius,
                             out.ctypes.data, out.size, C.byref(b))
    return out, b


def compute_normals(area, radius=1.5):
    L = lib()
    area = np.ascontiguousarray(area[:, :3], np.float32)
    out = np.zeros((area.shape[0], 3), np.float32)
    L.orc_compute_normals(area.ctypes.data, area.shape[0], radius, out.ctypes.data)
    return out


def generate_candidates(terrain: Index, bounds, fixed, num_candidates, search_radius, sensor_height):
    L = lib()
    fixed = np.ascontiguousarray(fixed, A.POSE_DTYPE).reshape(1)
    side = int(np.ceil(np.sqrt(num_candidates)))
    out = np.zeros(side * side, A.POSE_DTYPE)
    n = L.orc_generate_candidates(terrain.h, terrain.xyz.ctypes.data, C.byref(bounds), fixed.ctypes.data,
                                  num_candidates, search_radius, sensor_height, out.ctypes.data, out.size)
    return out[:n].copy()


def crop(xyz, box):
    L = lib()
    xyz = np.ascontiguousarray(xyz[:, :3], np.float32)
    out = np.zeros_like(xyz)
    idx = np.zeros(xyz.shape[0], np.int64)
    n = L.orc_crop(xyz.ctypes.data, xyz.shape[0], C.byref(box), out.ctypes.data, idx.ctypes.data)
    return out[:n].copy(), idx[:n].copy()


def voxel_grid(xyz, leaf):
    """Returns (centroids, keys, counts, point_keys) or None when PCL would return the input unchanged."""
    L = lib()
    xyz = np.ascontiguousarray(xyz[:, :3], np.float32)
    n = xyz.shape[0]
    out = np.zeros((max(n, 1), 3), np.float32)
    keys = np.zeros(max(n, 1), np.int32)
    counts = np.zeros(max(n, 1), np.int32)
    pk = np.zeros(max(n, 1), np.int32)
    v = L.orc_voxel_grid(xyz.ctypes.data, n, leaf, out.ctypes.data, keys.ctypes.data, counts.ctypes.data, pk.ctypes.data)
    if v < 0:
        return None
    return out[:v].copy(), keys[:v].copy(), counts[:v].copy(), pk[:n].copy()


def transform(xyz, t, q):
    L = lib()
    xyz = np.ascontiguousarray(xyz[:, :3], np.float32)
    t = np.ascontiguousarray(t, np.float64)
    q = np.ascontiguousarray(q, np.float64)
    out = np.zeros_like(xyz)
    L.orc_transform(xyz.ctypes.data, xyz.shape[0], t.ctypes.data, q.ctypes.data, out.ctypes.data)
    return out


def drivable_area(xyz, t, q, robot_xy, start_xy, *, grid_resolution=1.0, map_width=100.0, map_height=100.0,
                  max_gradient=0.3, min_points_per_cell=10, start_clear_radius=3.0):
    """calc_drivable_area.cpp:87-232.  Returns the (height, width) int8 grid, or None for an empty scan."""
    L = lib()
    xyz = np.ascontiguousarray(xyz[:, :3], np.float32)
    tp = None if t is None else np.ascontiguousarray(t, np.float64)
    qp = None if q is None else np.ascontiguousarray(q, np.float64)
    w, h = C.c_int32(0), C.c_int32(0)
    gw, gh = int(map_width / grid_resolution), int(map_height / grid_resolution)
    out = np.zeros(max(gw * gh, 1), np.int8)
    n = L.orc_drivable_area(xyz.ctypes.data, xyz.shape[0], A.ptr(tp), A.ptr(qp), robot_xy[0], robot_xy[1],
                            start_xy[0], start_xy[1], grid_resolution, map_width, map_height, max_gradient,
                            min_points_per_cell, start_clear_radius, out.ctypes.data, C.byref(w), C.byref(h))
    assert (w.value, h.value) == (gw, gh)
    if n < 0:
        return None
    return out[: gw * gh].reshape(gh, gw).copy()


def excavate(xyz, rgba, base_t, base_q, params, n_surface_bound: int, n_area_bound: int):
    """excavated_surface_generator.cpp:263-584.  Returns (terrain records (n, 32) uint8, area records, info)."""
    L = lib()
    xyz = np.ascontiguousarray(xyz[:, :3], np.float32)
    rg = None if rgba is None else np.ascontiguousarray(rgba, np.uint32)
    t = np.ascontiguousarray(base_t, np.float64)
    q = np.ascontiguousarray(base_q, np.float64)
    terrain = np.zeros((xyz.shape[0] + n_surface_bound, 32), np.uint8)
    area = np.zeros((max(n_area_bound, 1), 32), np.uint8)
    nt, na = C.c_uint64(0), C.c_uint64(0)
    info = A.ExcavationInfo()
    L.orc_excavate(xyz.ctypes.data, A.ptr(rg), xyz.shape[0], t.ctypes.data, q.ctypes.data, C.byref(params),
                   terrain.ctypes.data, C.byref(nt), area.ctypes.data, C.byref(na), C.byref(info))
    assert nt.value <= terrain.shape[0] and na.value <= area.shape[0]
    return terrain[: nt.value], area[: na.value], info
