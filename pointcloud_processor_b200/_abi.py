"""ctypes binding of the C ABI in include/plc.h and include/plc_host.h.

This is the headless stand-in for the cgo/JNI-style stub a maintainer would add to the reference's
ROS2 nodes (INTEGRATION.md shows the C++ call sites).  It loads the in-tree
``pointcloud_processor_b200/lib/libplc.so`` and fails loudly when it is missing: there is no
Python or CPU fallback for any plc_* entry point.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ["PLC_DEV_LIB"]) if os.environ.get("PLC_DEV_LIB") else _HERE / "lib" / "libplc.so"   # PLC_DEV_LIB: kernel experiments only

PLC_OK = 0
PLC_E_INVALID_ARG = -1
PLC_E_NO_TERRAIN = -2
PLC_E_NO_CELLS = -3
PLC_E_CAPACITY = -4
PLC_E_CUDA = -5
PLC_E_NCCL = -6
PLC_E_OOM = -7
PLC_E_NO_CANDIDATES = -8

FLAG_RANGE_FIXED, FLAG_FOV_FIXED, FLAG_VIS_FIXED = 1, 2, 4
FLAG_RANGE_MOBILE, FLAG_FOV_MOBILE, FLAG_VIS_MOBILE = 8, 16, 32


class CloudView(C.Structure):
    _fields_ = [("data", C.c_void_p), ("n_points", C.c_uint64), ("point_step", C.c_uint32),
                ("off_x", C.c_uint32), ("off_y", C.c_uint32), ("off_z", C.c_uint32)]


class Params(C.Structure):
    _fields_ = [("max_distance", C.c_double), ("min_distance", C.c_double), ("ray_step", C.c_double),
                ("ray_start", C.c_double), ("vis_radius", C.c_double), ("block_radius_scale", C.c_double),
                ("observed_radius_scale", C.c_double), ("fov_vertical", C.c_double),
                ("fov_horizontal", C.c_double), ("alpha", C.c_double), ("beta", C.c_double),
                ("fixed_mode", C.c_int32), ("reserved", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("rays_gated_in", C.c_uint64), ("rays_visible", C.c_uint64),
                ("samples", C.c_uint64), ("fixed_samples", C.c_uint64), ("exact_tests", C.c_uint64)]


class SweepOut(C.Structure):
    _fields_ = [("best_index", C.c_int64), ("best_score", C.c_double), ("best_covered", C.c_int32),
                ("n_cells", C.c_int32), ("fixed_total", C.c_double), ("counters", Counters),
                ("scores", C.c_void_p), ("covered", C.c_void_p), ("mask", C.c_void_p),
                ("mask_words", C.c_uint64), ("cell_flags", C.c_void_p), ("fixed_scores", C.c_void_p)]


class DeviceTimes(C.Structure):
    _fields_ = [("set_terrain_ms", C.c_float), ("set_observed_ms", C.c_float), ("filter_ms", C.c_float),
                ("merge_ms", C.c_float), ("sweep_kernel_ms", C.c_float), ("normals_ms", C.c_float),
                ("cells_ms", C.c_float), ("candidates_ms", C.c_float), ("drivable_ms", C.c_float),
                ("excavate_ms", C.c_float), ("sweep_fixed_ms", C.c_float), ("sweep_reduce_ms", C.c_float)]


class GridInfo(C.Structure):
    _fields_ = [("n_points", C.c_uint64), ("n_voxels", C.c_uint64), ("grid_bytes", C.c_uint64),
                ("dim", C.c_int32 * 3), ("origin", C.c_double * 3), ("voxel", C.c_double)]


class CropBox(C.Structure):
    _fields_ = [("x_min", C.c_double), ("x_max", C.c_double), ("y_min", C.c_double),
                ("y_max", C.c_double), ("z_min", C.c_double), ("z_max", C.c_double)]


class MergeSrc(C.Structure):
    _fields_ = [("cloud", CloudView), ("t", C.c_double * 3), ("q", C.c_double * 4),
                ("r", C.c_uint8), ("g", C.c_uint8), ("b", C.c_uint8), ("present", C.c_uint8)]


class GridBounds(C.Structure):
    _fields_ = [("min_x", C.c_double), ("max_x", C.c_double), ("min_y", C.c_double),
                ("max_y", C.c_double), ("min_z", C.c_double), ("max_z", C.c_double),
                ("width", C.c_int32), ("height", C.c_int32)]


class ColorStats(C.Structure):
    _fields_ = [("total", C.c_int32), ("green", C.c_int32), ("red", C.c_int32), ("blue", C.c_int32),
                ("yellow", C.c_int32)]


class Pose(C.Structure):
    _fields_ = [("x", C.c_double), ("y", C.c_double), ("z", C.c_double), ("pitch", C.c_double),
                ("yaw", C.c_double)]


class NodeResult(C.Structure):      # plc_vlnode_result (include/plc_node.h)
    _fields_ = [("ran", C.c_int32), ("n_candidates", C.c_int32), ("best", Pose), ("best_score", C.c_double),
                ("fixed", Pose), ("fixed_total", C.c_double), ("range_ok", C.c_int32), ("fov_ok", C.c_int32),
                ("visibility_ok", C.c_int32), ("n_cells", C.c_int32), ("zx120_only", ColorStats),
                ("dual", ColorStats), ("observed_points", C.c_uint64)]


class MergerInput(C.Structure):     # plc_merger_input
    _fields_ = [("cloud", CloudView), ("has_cloud", C.c_int32), ("has_transform", C.c_int32),
                ("t", C.c_double * 3), ("q", C.c_double * 4)]


class ExcavationParams(C.Structure):  # plc_excavation_params
    _fields_ = [("depth", C.c_double), ("slope_angle_deg", C.c_double), ("offset_x", C.c_double), ("offset_y", C.c_double),
                ("point_density", C.c_double), ("terrain_search_radius", C.c_double),
                ("arm1_length", C.c_double), ("arm1_width", C.c_double), ("arm2_length", C.c_double),
                ("arm2_width", C.c_double), ("width", C.c_double), ("length", C.c_double),
                ("l_shape_enabled", C.c_int32), ("area_walls_vertical", C.c_int32)]


class ExcavationInfo(C.Structure):    # plc_excavation_info
    _fields_ = [("center", C.c_double * 3), ("yaw", C.c_double), ("n_kept", C.c_uint64), ("n_surface", C.c_uint64),
                ("n_area", C.c_uint64)]


class DrivableParams(C.Structure):  # plc_drivable_params
    _fields_ = [("grid_resolution", C.c_double), ("map_width", C.c_double), ("map_height", C.c_double),
                ("max_gradient", C.c_double), ("start_clear_radius", C.c_double),
                ("min_points_per_cell", C.c_int32), ("reserved", C.c_int32)]


class DrivableInfo(C.Structure):    # plc_drivable_info
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("origin_x", C.c_double), ("origin_y", C.c_double),
                ("resolution", C.c_double), ("n_cells", C.c_uint64), ("published", C.c_int32), ("reserved", C.c_int32)]


# numpy mirrors of the PODs (same byte layout as plc_cell / plc_pose)
CELL_DTYPE = np.dtype([("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("nx", "<f4"), ("ny", "<f4"),
                       ("nz", "<f4"), ("pad", "<u4")])
POSE_DTYPE = np.dtype([("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("pitch", "<f8"), ("yaw", "<f8")])
assert CELL_DTYPE.itemsize == 40 and POSE_DTYPE.itemsize == 40

# every symbol include/plc.h, plc_host.h and plc_node.h declare (tests check the library exports them all)
EXPORTS = [
    "plc_params_default", "plc_params_memo", "plc_ctx_create", "plc_ctx_destroy", "plc_strerror",
    "plc_last_error", "plc_launch_count", "plc_abi_version", "plc_ctx_sync",
    "plc_vl_set_terrain", "plc_vl_set_observed", "plc_vl_set_cells", "plc_vl_sweep",
    "plc_vl_upload_candidates", "plc_vl_run_async", "plc_vl_fetch", "plc_vl_result_device_ptr",
    "plc_vl_grid_info", "plc_vl_last_sweep_kernel_ms", "plc_last_device_times", "plc_vl_count_samples", "plc_vl_graph_replays",
    "plc_vl_upload_candidates_strided", "plc_vl_sweep_strided", "plc_filter", "plc_merge",
    "plc_vl_compute_normals", "plc_vl_generate_cells", "plc_vl_generate_candidates",
    "plc_drivable_params_default", "plc_drivable_area", "plc_drivable_area_static",
    "plc_dev_alloc", "plc_dev_free", "plc_dev_upload", "plc_dev_download", "plc_dev_export", "plc_dev_import",
    "plc_dev_release",
    "plc_excavation_params_default", "plc_excavation_counts", "plc_excavate",
    "plc_host_generate_cells", "plc_host_compute_normals", "plc_host_generate_candidates",
    "plc_vlnode_create", "plc_vlnode_create_on_group", "plc_vlnode_destroy", "plc_vlnode_set_parameter_double", "plc_vlnode_set_parameter_int", "plc_vlnode_update_parameters",
    "plc_vlnode_excavation_area_callback", "plc_vlnode_terrain_callback", "plc_vlnode_zx120_points_callback",
    "plc_vlnode_set_base_transform", "plc_vlnode_run_optimization", "plc_vlnode_get_candidates",
    "plc_vlnode_get_cells", "plc_vlnode_get_bounds", "plc_vlnode_take_log", "plc_filter_scan", "plc_merge_clouds",
    "plc_group_create", "plc_group_destroy", "plc_group_size", "plc_group_ctx", "plc_group_last_error",
    "plc_group_set_terrain", "plc_group_set_observed", "plc_group_set_cells", "plc_group_sweep",
    "plc_vl_set_terrain_unindexed", "plc_group_set_terrain_unindexed",
]

_lib = None


class PlcError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str = ""):
        self.status = status
        msg = f"{where}: {strerror(status)} ({status})"
        if detail:
            msg += f" [{detail}]"
        super().__init__(msg)


def lib() -> C.CDLL:
    """Loads libplc.so (once).  Missing library is an error, never a fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C pointcloud_processor_b200/csrc`. There is no CPU fallback.")
    L = C.CDLL(os.fspath(LIB_PATH))
    vp, u64, i32, dbl = C.c_void_p, C.c_uint64, C.c_int32, C.c_double
    L.plc_strerror.restype = C.c_char_p
    L.plc_strerror.argtypes = [C.c_int]
    L.plc_last_error.restype = C.c_char_p
    L.plc_last_error.argtypes = [vp]
    L.plc_launch_count.restype = u64
    L.plc_launch_count.argtypes = [vp]
    L.plc_abi_version.restype = C.c_int
    L.plc_params_default.argtypes = [C.POINTER(Params)]
    L.plc_params_default.restype = None
    L.plc_params_memo.argtypes = [C.POINTER(Params)]
    L.plc_params_memo.restype = None
    L.plc_ctx_create.argtypes = [C.c_int, vp, C.POINTER(vp)]
    L.plc_ctx_destroy.argtypes = [vp]
    L.plc_ctx_destroy.restype = None
    L.plc_ctx_sync.argtypes = [vp]
    for name in ("plc_vl_set_terrain", "plc_vl_set_observed"):
        getattr(L, name).argtypes = [vp, C.POINTER(CloudView)]
    L.plc_vl_set_cells.argtypes = [vp, vp, u64]
    L.plc_vl_sweep.argtypes = [vp, C.POINTER(Params), vp, vp, u64, u64, u64, C.POINTER(SweepOut)]
    L.plc_vl_upload_candidates.argtypes = [vp, C.POINTER(Params), vp, vp, u64, u64, u64]
    L.plc_vl_run_async.argtypes = [vp, C.c_int, C.c_int]
    L.plc_vl_fetch.argtypes = [vp, C.POINTER(SweepOut)]
    L.plc_vl_result_device_ptr.argtypes = [vp, C.POINTER(vp)]
    L.plc_vl_last_sweep_kernel_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.plc_last_device_times.argtypes = [vp, C.POINTER(DeviceTimes)]
    L.plc_vl_upload_candidates_strided.argtypes = [vp, C.POINTER(Params), vp, vp, u64, u64, u64]
    L.plc_vl_sweep_strided.argtypes = [vp, C.POINTER(Params), vp, vp, u64, u64, u64, C.POINTER(SweepOut)]
    L.plc_vl_grid_info.argtypes = [vp, C.POINTER(GridInfo), C.POINTER(GridInfo)]
    L.plc_filter.argtypes = [vp, C.POINTER(CloudView), C.POINTER(CropBox), dbl, vp, u64,
                             C.POINTER(u64), C.POINTER(u64)]
    L.plc_merge.argtypes = [vp, C.POINTER(MergeSrc), C.c_int, vp, u64, C.POINTER(u64),
                            C.POINTER(vp), C.POINTER(u64)]
    L.plc_host_generate_cells.argtypes = [C.POINTER(CloudView), vp, dbl, i32, dbl, vp, u64,
                                          C.POINTER(u64), C.POINTER(GridBounds)]
    L.plc_host_compute_normals.argtypes = [C.POINTER(CloudView), dbl, vp]
    L.plc_vl_generate_cells.argtypes = [vp, C.POINTER(CloudView), vp, dbl, i32, dbl, vp, u64,
                                        C.POINTER(u64), C.POINTER(GridBounds)]
    L.plc_vl_compute_normals.argtypes = [vp, C.POINTER(CloudView), dbl, vp]
    L.plc_dev_alloc.argtypes = [vp, u64, C.POINTER(vp)]
    L.plc_dev_free.argtypes = [vp, vp]
    L.plc_dev_free.restype = None
    L.plc_dev_upload.argtypes = [vp, vp, vp, u64]
    L.plc_dev_download.argtypes = [vp, vp, vp, u64]
    L.plc_dev_export.argtypes = [vp, vp, vp]
    L.plc_dev_import.argtypes = [vp, vp, C.POINTER(vp)]
    L.plc_dev_release.argtypes = [vp, vp]
    L.plc_dev_release.restype = None
    L.plc_excavation_params_default.argtypes = [C.POINTER(ExcavationParams)]
    L.plc_excavation_params_default.restype = None
    L.plc_excavation_counts.argtypes = [C.POINTER(ExcavationParams), C.POINTER(u64), C.POINTER(u64)]
    L.plc_excavate.argtypes = [vp, C.POINTER(CloudView), C.c_uint32, vp, vp, C.POINTER(ExcavationParams), vp, u64,
                               C.POINTER(u64), vp, u64, C.POINTER(u64), C.POINTER(ExcavationInfo)]
    L.plc_drivable_params_default.argtypes = [C.POINTER(DrivableParams)]
    L.plc_drivable_params_default.restype = None
    L.plc_drivable_area.argtypes = [vp, C.POINTER(CloudView), vp, vp, dbl, dbl, dbl, dbl, C.POINTER(DrivableParams),
                                    vp, u64, C.POINTER(DrivableInfo)]
    L.plc_drivable_area_static.argtypes = L.plc_drivable_area.argtypes
    L.plc_vl_count_samples.argtypes = [vp, C.c_int]
    L.plc_vl_graph_replays.argtypes = [vp, C.POINTER(u64)]
    L.plc_vl_generate_candidates.argtypes = [vp, C.POINTER(CloudView), C.POINTER(GridBounds), vp, i32,
                                             dbl, dbl, vp, u64, C.POINTER(u64)]
    L.plc_host_generate_candidates.argtypes = [C.POINTER(CloudView), C.POINTER(GridBounds), vp, i32,
                                               dbl, dbl, vp, u64, C.POINTER(u64)]
    L.plc_vlnode_create.argtypes = [vp, C.POINTER(vp)]
    L.plc_vlnode_create_on_group.argtypes = [vp, C.POINTER(vp)]
    L.plc_vlnode_destroy.argtypes = [vp]
    L.plc_vlnode_destroy.restype = None
    L.plc_vlnode_set_parameter_double.argtypes = [vp, C.c_char_p, dbl]
    L.plc_vlnode_set_parameter_int.argtypes = [vp, C.c_char_p, C.c_int64]
    L.plc_vlnode_update_parameters.argtypes = [vp]
    for name in ("plc_vlnode_excavation_area_callback", "plc_vlnode_terrain_callback",
                 "plc_vlnode_zx120_points_callback"):
        getattr(L, name).argtypes = [vp, C.POINTER(CloudView)]
    L.plc_vlnode_set_base_transform.argtypes = [vp, i32, dbl, dbl, dbl]
    L.plc_vlnode_run_optimization.argtypes = [vp, C.POINTER(NodeResult)]
    L.plc_vlnode_get_candidates.argtypes = [vp, vp, u64, C.POINTER(u64)]
    L.plc_vlnode_get_cells.argtypes = [vp, vp, vp, u64, C.POINTER(u64)]
    L.plc_vlnode_get_bounds.argtypes = [vp, C.POINTER(GridBounds)]
    L.plc_vlnode_take_log.argtypes = [vp]
    L.plc_vlnode_take_log.restype = C.c_char_p
    L.plc_filter_scan.argtypes = [vp, C.POINTER(CloudView), dbl, dbl, dbl, dbl, vp, u64, C.POINTER(u64)]
    L.plc_merge_clouds.argtypes = [vp, C.POINTER(MergerInput), C.POINTER(MergerInput), vp, u64, C.POINTER(u64),
                                   vp, u64, C.POINTER(u64), vp, u64, C.POINTER(u64)]
    L.plc_group_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
    L.plc_group_destroy.argtypes = [vp]
    L.plc_group_destroy.restype = None
    L.plc_group_size.argtypes = [vp]
    L.plc_group_ctx.argtypes = [vp, C.c_int]
    L.plc_group_ctx.restype = vp
    L.plc_group_last_error.argtypes = [vp]
    L.plc_group_last_error.restype = C.c_char_p
    L.plc_group_set_terrain.argtypes = [vp, C.POINTER(CloudView)]
    L.plc_group_set_observed.argtypes = [vp, C.POINTER(CloudView)]
    L.plc_group_set_cells.argtypes = [vp, vp, u64]
    L.plc_group_sweep.argtypes = [vp, C.POINTER(Params), vp, vp, u64, C.POINTER(SweepOut)]
    _lib = L
    return L


def strerror(status: int) -> str:
    return lib().plc_strerror(int(status)).decode()


def default_params() -> Params:
    p = Params()
    lib().plc_params_default(C.byref(p))
    return p


def memo_params() -> Params:
    p = Params()
    lib().plc_params_memo(C.byref(p))
    return p


def excavation_params(**kw) -> ExcavationParams:
    p = ExcavationParams()
    lib().plc_excavation_params_default(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def excavation_counts(p: ExcavationParams):
    ns, na = C.c_uint64(0), C.c_uint64(0)
    st = lib().plc_excavation_counts(C.byref(p), C.byref(ns), C.byref(na))
    if st != PLC_OK:
        raise PlcError(st, "plc_excavation_counts")
    return ns.value, na.value


def cloud_view(buf: np.ndarray, point_step: int | None = None, off=(0, 4, 8)) -> CloudView:
    """View over a PointCloud2-style buffer.

    ``buf`` is either an (n, k) float32 array whose first three columns are x, y, z (point_step =
    4k) or a flat uint8 array with an explicit ``point_step``.  The caller keeps ``buf`` alive.
    """
    if isinstance(buf, CloudView):          # already a view (e.g. DevBuffer.view(): device memory)
        return buf
    if buf.dtype == np.uint8 and buf.ndim == 1:
        if not point_step:
            raise ValueError("point_step required for a byte buffer")
        n = buf.size // point_step
    else:
        if buf.dtype != np.float32 or buf.ndim != 2 or buf.shape[1] < 3:
            raise ValueError("expected an (n, >=3) float32 array or a uint8 byte buffer")
        point_step = buf.shape[1] * 4
        n = buf.shape[0]
    if not buf.flags["C_CONTIGUOUS"]:
        raise ValueError("cloud buffer must be C-contiguous")
    v = CloudView()
    v.data = buf.ctypes.data if n else None
    v.n_points = n
    v.point_step = point_step
    v.off_x, v.off_y, v.off_z = off
    return v


def ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data
