"""SURVEY §8 N3 across processes: a cloud one process left in a device buffer is the terrain / filter input of another
(upstream: pointcloud_filter -> pointcloud_merger -> excavated_surface_generator -> virtual_lidar are separate ROS nodes,
launch/optimal_viewpoint.launch.py).  The handle travels as 64 opaque bytes; the bytes of the cloud never leave HBM."""
import json
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent

CONSUMER = r"""
import json, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from pointcloud_processor_b200 import _abi as A
from pointcloud_processor_b200.context import Context, ImportedBuffer
handle, n, leaf = bytes.fromhex(sys.argv[2]), int(sys.argv[3]), float(sys.argv[4])
with Context(0) as ctx:
    buf = ImportedBuffer(ctx, handle, n * 16)
    raw = buf.download(n * 16, np.float32).reshape(n, 4)
    box = A.CropBox(-50.0, 50.0, -50.0, 50.0, -50.0, 50.0)
    v = buf.view(n, 16)
    out = np.zeros((n, 4), np.float32)
    cnt, n_crop = A.C.c_uint64(0), A.C.c_uint64(0)
    st = ctx._L.plc_filter(ctx._h, A.C.byref(v), A.C.byref(box), leaf, out.ctypes.data, n, A.C.byref(cnt), A.C.byref(n_crop))
    assert st == A.PLC_OK, st
    buf.free()
print(json.dumps({"checksum": float(raw.astype(np.float64).sum()), "n_filtered": int(cnt.value),
                  "filtered_sum": float(out[: cnt.value, :3].astype(np.float64).sum())}))
"""


def test_exported_buffer_is_another_process_input(tmp_path):
    from pointcloud_processor_b200 import _abi as A
    from pointcloud_processor_b200.context import Context, DevBuffer, ImportedBuffer

    rng = np.random.default_rng(7)
    n, leaf = 50_000, 0.2
    cloud = np.zeros((n, 4), np.float32)
    cloud[:, :3] = rng.uniform(-10, 10, (n, 3)).astype(np.float32)
    with Context(0) as ctx:
        buf = DevBuffer(ctx, cloud.nbytes).upload(cloud)
        handle = buf.export()
        assert len(handle) == 64
        # the exporting process cannot map its own handle: a plain error, not a crash
        with pytest.raises(A.PlcError):
            ImportedBuffer(ctx, handle, cloud.nbytes)
        script = tmp_path / "consumer.py"
        script.write_text(CONSUMER)
        r = subprocess.run([sys.executable, str(script), str(ROOT), handle.hex(), str(n), str(leaf)],
                           capture_output=True, text=True, timeout=300)
        if r.returncode != 0 and ("not supported" in r.stderr or "operation not permitted" in r.stderr.lower()):
            pytest.skip("CUDA IPC is not available between processes on this box: " + r.stderr.strip().splitlines()[-1])
        assert r.returncode == 0, r.stderr[-2000:]
        got = json.loads(r.stdout.strip().splitlines()[-1])
        # the same call in this process on the same device buffer
        ref, _ = ctx.filter(cloud, A.CropBox(-50.0, 50.0, -50.0, 50.0, -50.0, 50.0), leaf)
        assert got["checksum"] == float(cloud.astype(np.float64).sum())
        assert got["n_filtered"] == ref.shape[0]
        assert got["filtered_sum"] == float(np.asarray(ref)[:, :3].astype(np.float64).sum())
        # the producer still owns the memory after the consumer has let go of it
        assert np.array_equal(buf.download(cloud.nbytes, np.float32).reshape(n, 4), cloud)
        buf.free()


def test_import_of_garbage_is_an_error_not_a_crash():
    from pointcloud_processor_b200 import _abi as A
    from pointcloud_processor_b200.context import Context, ImportedBuffer

    with Context(0) as ctx:
        for handle in (bytes(64), bytes(range(64))):
            with pytest.raises(A.PlcError):
                ImportedBuffer(ctx, handle, 1024)
        # the context is still usable afterwards (no sticky error)
        cloud = np.zeros((100, 4), np.float32)
        cloud[:, 0] = np.arange(100, dtype=np.float32) * 0.01
        out, n_crop = ctx.filter(cloud, A.CropBox(-1.0, 2.0, -1.0, 1.0, -1.0, 1.0), 0.05)
        assert n_crop == 100 and 0 < out.shape[0] <= 100
