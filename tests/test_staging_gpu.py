"""Large pageable inputs go to the device through the library's piecewise staging (plc_ctx.cu upload_pageable): the
bytes that arrive are the bytes that were given, at piece boundaries and off them, call after call."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nbytes", [(32 << 20) - 1, 32 << 20, (32 << 20) + 1, (33 << 20) + 12345, (70 << 20) + 17])
def test_pageable_upload_round_trip(nbytes):
    from pointcloud_processor_b200.context import Context, DevBuffer

    rng = np.random.default_rng(nbytes)
    with Context(0) as ctx:
        buf = DevBuffer(ctx, nbytes)
        for rep in range(3):                       # the staging buffer is reused: a later call must not race an earlier one
            data = rng.integers(0, 256, nbytes, dtype=np.uint8)
            buf.upload(data)
            assert np.array_equal(buf.download(nbytes), data)
        buf.free()


def test_filter_of_a_large_pageable_scan_equals_the_device_resident_one():
    from pointcloud_processor_b200 import _abi as A
    from pointcloud_processor_b200.context import Context, DevBuffer

    rng = np.random.default_rng(11)
    n = 2_500_000
    scan = np.zeros((n, 4), np.float32)
    scan[:, :3] = rng.uniform(-20, 20, (n, 3)).astype(np.float32)
    box = A.CropBox(-15.0, 15.0, -10.0, 10.0, -10.0, 10.0)
    with Context(0) as ctx:
        a, crop_a = ctx.filter(scan, box, 0.2)                       # 40 MB of pageable memory: staged piecewise
        b, crop_b = ctx.filter(scan[: n // 2], box, 0.2)             # a second, smaller call (plain copy) right after
        buf = DevBuffer(ctx, scan.nbytes).upload(scan)
        v = buf.view(n, 16)
        out = np.zeros((n, 4), np.float32)
        n_out, n_crop = A.C.c_uint64(0), A.C.c_uint64(0)
        st = ctx._L.plc_filter(ctx._h, A.C.byref(v), A.C.byref(box), 0.2, out.ctypes.data, n, A.C.byref(n_out), A.C.byref(n_crop))
        assert st == A.PLC_OK
        assert crop_a == n_crop.value and np.array_equal(a, out[: n_out.value])
        v2 = buf.view(n // 2, 16)
        st = ctx._L.plc_filter(ctx._h, A.C.byref(v2), A.C.byref(box), 0.2, out.ctypes.data, n, A.C.byref(n_out), A.C.byref(n_crop))
        assert st == A.PLC_OK
        assert crop_b == n_crop.value and np.array_equal(b, out[: n_out.value])
        buf.free()
