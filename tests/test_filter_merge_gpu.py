"""GPU parity tests of the pointcloud_filter (crop + VoxelGrid) and pointcloud_merger (rigid
transform + colour + concat) kernels against the oracle.

Bar (north_star): bit-exact in index and voxel membership, coordinates within 1e-6 relative.  The
kernels sum each voxel in (voxel index, point index) order, the same order the oracle fixes, so the
tests additionally expect identical float32 bits."""
import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import _abi as A

pytestmark = pytest.mark.gpu

COORD_RTOL = 1e-6
DEFAULT_BOX = A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0)     # pointcloud_filter.cpp:30-36, :111-113


def scan(n, seed, step_floats=8):
    """Synthetic 32-ring scan with 2 cm range noise, PointCloud2 record of step_floats*4 bytes."""
    rng = np.random.default_rng(seed)
    az = rng.uniform(-np.pi, np.pi, n)
    el = np.deg2rad(rng.integers(0, 32, n) * (40.0 / 31) - 25.0)
    rng_m = np.minimum(60.0, 2.2 / np.maximum(np.sin(-el), 0.02)) + rng.normal(0, 0.02, n)
    pts = np.zeros((n, step_floats), np.float32)
    pts[:, 0] = rng_m * np.cos(el) * np.cos(az)
    pts[:, 1] = rng_m * np.cos(el) * np.sin(az)
    pts[:, 2] = rng_m * np.sin(el) + 2.0
    if step_floats > 3:
        pts[:, 3] = rng.uniform(0, 255, n)      # intensity: ignored by fromROSMsg field mapping
    return pts


def check_filter(ctx, pts, box, leaf):
    out, n_crop = ctx.filter(pts, box, leaf)
    cxyz, _ = O.crop(pts, box)
    assert n_crop == cxyz.shape[0]
    if leaf <= 0 or cxyz.shape[0] == 0:
        assert out.shape[0] == cxyz.shape[0]
        assert np.array_equal(out[:, :3].view(np.uint32), cxyz.view(np.uint32))
        return out
    vg = O.voxel_grid(cxyz, leaf)
    if vg is None:                                    # PCL: leaf too small => input returned unchanged
        assert np.array_equal(out[:, :3].view(np.uint32), cxyz.view(np.uint32))
        return out
    cent, keys, counts, _ = vg
    assert out.shape[0] == cent.shape[0]              # same voxel membership => same voxel count
    assert np.all(np.diff(keys) > 0)                  # ascending voxel index, one output per voxel
    np.testing.assert_allclose(out[:, :3], cent, rtol=COORD_RTOL, atol=1e-6)
    assert np.array_equal(out[:, :3].view(np.uint32), cent.view(np.uint32))
    assert np.all(out[:, 3] == 1.0)                   # pcl::PointXYZ data[3]
    return out


def test_filter_default_box(ctx):
    check_filter(ctx, scan(400_000, 4), DEFAULT_BOX, 0.2)


def test_filter_crop_only_and_strict_bounds(ctx):
    pts = scan(50_000, 5, step_floats=4)
    pts[:6, :3] = [[0.0, 0, 0], [15.0, 0, 0], [5, -10.0, 0], [5, 10.0, 0], [5, 0, -1.5], [5, 0, 10.0]]  # on a face: out
    pts[6:9, :3] = [[np.nan, 0, 0], [1, np.inf, 0], [1, 1, np.nan]]                                     # NaN fails
    pts[9, :3] = [np.nextafter(np.float32(0), np.float32(1)), 0, 0]                                      # just inside
    out = check_filter(ctx, pts, DEFAULT_BOX, 0.0)
    assert out.shape[0] > 0


def test_filter_leaf_sizes_and_layouts(ctx):
    pts = scan(120_000, 6, step_floats=3)
    for leaf in (0.05, 0.2, 0.5, 3.0):
        check_filter(ctx, pts, DEFAULT_BOX, leaf)
    # unaligned 22-byte velodyne-style record
    raw = np.zeros((pts.shape[0], 22), np.uint8)
    raw[:, 2:14] = pts.view(np.uint8).reshape(-1, 12)
    v = ctx.filter(raw.reshape(-1), DEFAULT_BOX, 0.2, point_step=22, off=(2, 6, 10))[0]
    w = ctx.filter(pts, DEFAULT_BOX, 0.2)[0]
    assert np.array_equal(v, w)


def test_filter_empty_and_degenerate(ctx):
    out, n = ctx.filter(np.zeros((0, 3), np.float32), DEFAULT_BOX, 0.2)
    assert out.shape[0] == 0 and n == 0
    far = scan(1000, 7, 3) + np.float32(500.0)          # nothing inside the box: empty cloud is published (Q7)
    out, n = ctx.filter(far, DEFAULT_BOX, 0.2)
    assert out.shape[0] == 0 and n == 0
    one = np.array([[1.0, 1.0, 1.0]], np.float32)
    out = check_filter(ctx, one, DEFAULT_BOX, 0.2)
    assert out.shape[0] == 1
    # leaf so small that the voxel count overflows int32: PCL returns the cropped input unchanged
    big = A.CropBox(-1e4, 1e4, -1e4, 1e4, -1e4, 1e4)
    wide = (np.random.default_rng(8).uniform(-900, 900, (5000, 3))).astype(np.float32)
    out = check_filter(ctx, wide, big, 0.001)
    assert out.shape[0] == 5000


def quat_rpy(roll, pitch, yaw):
    cr, sr, cp, sp, cy, sy = np.cos(roll / 2), np.sin(roll / 2), np.cos(pitch / 2), np.sin(pitch / 2), np.cos(yaw / 2), np.sin(yaw / 2)
    return [sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy]


def test_merge_two_sources(ctx):
    robot = ctx.filter(scan(200_000, 9), DEFAULT_BOX, 0.2)[0]
    backhoe = ctx.filter(scan(200_000, 10), DEFAULT_BOX, 0.2)[0]
    # static TFs of launch/tf_*.launch.xml composed with a base pose
    srcs = [dict(cloud=robot, t=[12.25, -3.0, 2.0], q=quat_rpy(0, 0.1245, 0.7), rgb=(255, 0, 0)),
            dict(cloud=backhoe, t=[0.55, 0.4, 3.5], q=quat_rpy(0, 0.4363, 0), rgb=(0, 0, 255))]
    merged, each = ctx.merge(srcs)
    exp = [O.transform(s["cloud"], s["t"], s["q"]) for s in srcs]
    n0, n1 = exp[0].shape[0], exp[1].shape[0]
    assert merged.shape == (n0 + n1, 32) and [e.shape[0] for e in each] == [n0, n1]
    f = merged.view(np.float32).reshape(-1, 8)
    u = merged.view(np.uint32).reshape(-1, 8)
    np.testing.assert_allclose(f[:, :3], np.concatenate(exp), rtol=COORD_RTOL, atol=1e-6)
    assert np.array_equal(f[:, :3].view(np.uint32), np.concatenate(exp).view(np.uint32))
    assert np.all(f[:, 3] == 1.0) and np.all(u[:, 5:] == 0)
    assert np.all(u[:n0, 4] == 0xFFFF0000) and np.all(u[n0:, 4] == 0xFF0000FF)   # robot red first, then backhoe blue
    assert np.array_equal(each[0], merged[:n0]) and np.array_equal(each[1], merged[n0:])


def test_merge_missing_source_and_empty(ctx):
    pts = scan(1000, 11, 4)
    srcs = [dict(cloud=pts, t=[0, 0, 0], q=[0, 0, 0, 1], rgb=(255, 0, 0), present=False),   # TF lookup failed: skipped
            dict(cloud=pts, t=[1, 2, 3], q=[0, 0, 0, 1], rgb=(0, 0, 255))]
    merged, each = ctx.merge(srcs)
    assert merged.shape[0] == 1000 and each[0].shape[0] == 0 and each[1].shape[0] == 1000
    f = merged.view(np.float32).reshape(-1, 8)
    assert np.array_equal(f[:, :3], pts[:, :3] + np.array([1, 2, 3], np.float32))
    merged, each = ctx.merge([dict(cloud=np.zeros((0, 3), np.float32), t=[0, 0, 0], q=[0, 0, 0, 1], rgb=(1, 2, 3))])
    assert merged.shape[0] == 0
