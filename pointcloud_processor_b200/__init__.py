"""pointcloud_processor_b200 — B200-native (sm_100a) implementation of the data-parallel hot path of
YamaguchiAtsushi/pointcloud_processor: the virtual_lidar visibility sweep and its filter/merge
feeders, behind the C ABI in include/plc.h.  The CUDA library is the product; Python here is only
the headless driver (ctypes) used by tests and bench.py."""
from . import _abi
from ._abi import PlcError, default_params, memo_params
from .context import Context, SweepResult, host_compute_normals, host_generate_candidates, host_generate_cells

__all__ = ["_abi", "PlcError", "default_params", "memo_params", "Context", "SweepResult",
           "host_compute_normals", "host_generate_candidates", "host_generate_cells"]
