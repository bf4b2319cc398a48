"""Independent NumPy restatements of the feeder stages, checked against oracle/ (CPU only).

The sweep has its SciPy twin in test_oracle_independent.py; this file does the same for the stages around it, so that
every oracle function the GPU parity tests lean on is pinned by a second implementation that shares no code with
oracle/*.cpp or with the CUDA path (it is the check on the third-party arithmetic that the reference-source fixtures of
test_reference_source.py leave restated).

Restated from /root/reference (arithmetic of the third-party calls as listed in SURVEY section A.3):
  src/pointcloud_filter.cpp:106-116   crop: strict six-sided box on the float promoted to double
  src/pointcloud_filter.cpp:129-136   pcl::VoxelGrid<PointXYZ>: float32 inverse leaf, floor(p * inv) - min_b, index
                                      i + j dx + k dx dy, one centroid per index in ascending order, float32 running sum
  src/pointcloud_merger.cpp:369-387   tf2::doTransform: Eigen quaternion -> rotation and p' = R p + t, all in float32
  src/calc_drivable_area.cpp:141-232  binning by static_cast<int>((p - origin) / res), gradient (max z - min z) / res
NumPy evaluates every float32 operation with one rounding and never fuses a multiply with an add, which is the
arithmetic of the reference's x86-64 build."""
import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import _abi as A

f32 = np.float32


def scan(n, seed, spread=22.0):
    rng = np.random.default_rng(seed)
    p = rng.uniform(-spread, spread, (n, 3)).astype(f32)
    p[:, 2] = (rng.uniform(-3.0, 12.0, n)).astype(f32)
    p[rng.choice(n, 40, replace=False), rng.integers(0, 3, 40)] = np.nan
    p[rng.choice(n, 10, replace=False), 0] = np.inf
    p[:6] = [[0.0, 1.0, 1.0], [15.0, 1.0, 1.0], [3.0, 10.0, 1.0], [3.0, -10.0, 1.0], [3.0, 1.0, -1.5], [3.0, 1.0, 10.0]]   # on the faces: out
    return p


def numpy_crop(p, box):
    d = p.astype(np.float64)
    keep = ((d[:, 0] > box.x_min) & (d[:, 0] < box.x_max) & (d[:, 1] > box.y_min) & (d[:, 1] < box.y_max) &
            (d[:, 2] > box.z_min) & (d[:, 2] < box.z_max))            # NaN fails every comparison
    return p[keep], np.flatnonzero(keep)


def numpy_voxel_grid(p, leaf):
    inv = f32(1.0) / f32(leaf)
    mn, mx = p.min(axis=0), p.max(axis=0)                              # getMinMax3D over the (finite) cloud
    min_b = np.floor(mn * inv).astype(np.int32)
    max_b = np.floor(mx * inv).astype(np.int32)
    div = (max_b - min_b + 1).astype(np.int64)
    ijk = (np.floor(p * inv) - min_b.astype(f32)).astype(np.int32)     # float32 subtraction, then the int cast
    idx = ijk[:, 0].astype(np.int64) + ijk[:, 1] * div[0] + ijk[:, 2] * div[0] * div[1]
    order = np.argsort(idx, kind="stable")                             # ties in point order (what the oracle fixes)
    heads = np.flatnonzero(np.r_[True, idx[order][1:] != idx[order][:-1]])
    ends = np.r_[heads[1:], idx.size]
    out = np.zeros((heads.size, 3), f32)
    for v, (a, b) in enumerate(zip(heads, ends)):
        acc = np.zeros(3, f32)
        for q in p[order[a:b]]:
            acc = acc + q                                              # float32 running sum, one point at a time
        out[v] = acc / f32(b - a)
    return out, idx[order][heads]


def numpy_rotation(q):
    x, y, z, w = (f32(v) for v in q)                                   # Eigen::Quaternionf::toRotationMatrix
    tx, ty, tz = f32(2) * x, f32(2) * y, f32(2) * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    one = f32(1)
    return np.array([[one - (tyy + tzz), txy - twz, txz + twy],
                     [txy + twz, one - (txx + tzz), tyz - twx],
                     [txz - twy, tyz + twx, one - (txx + tyy)]], f32)


def numpy_transform(p, t, q):
    R, tf = numpy_rotation(q), np.asarray(t, np.float64).astype(f32)
    out = np.empty_like(p)
    with np.errstate(invalid="ignore"):                                # non-finite points travel through, as upstream
        for k in range(3):
            acc = R[k, 0] * p[:, 0]
            acc = acc + R[k, 1] * p[:, 1]
            acc = acc + R[k, 2] * p[:, 2]
            out[:, k] = acc + tf[k]
    return out


def numpy_drivable(p, t, q, robot, start, res, width, height, max_gradient, min_points, clear_radius):
    m = numpy_transform(p, t, q) if t is not None else p
    m = m[np.all(np.isfinite(m), axis=1)]
    w, h = int(width / res), int(height / res)
    ox, oy = robot[0] - width / 2.0, robot[1] - height / 2.0
    with np.errstate(invalid="ignore"):
        gx = np.trunc((m[:, 0].astype(np.float64) - ox) / res)         # static_cast<int> truncates towards zero
        gy = np.trunc((m[:, 1].astype(np.float64) - oy) / res)
    ok = (gx >= 0) & (gx < w) & (gy >= 0) & (gy < h)
    cell = (gy[ok] * w + gx[ok]).astype(np.int64)
    z = m[ok, 2]
    count = np.bincount(cell, minlength=w * h)
    zmin = np.full(w * h, np.inf, f32)
    zmax = np.full(w * h, -np.inf, f32)
    np.minimum.at(zmin, cell, z)
    np.maximum.at(zmax, cell, z)
    yy, xx = np.divmod(np.arange(w * h), w)
    cx, cy = ox + (xx + 0.5) * res, oy + (yy + 0.5) * res
    dist = np.sqrt((cx - start[0]) ** 2 + (cy - start[1]) ** 2)
    with np.errstate(invalid="ignore"):
        gradient = np.where(count >= 2, ((zmax - zmin).astype(f32).astype(np.float64) / res).astype(f32), f32(0))
    grid = np.where(gradient.astype(np.float64) > max_gradient, 100, 0).astype(np.int8)
    grid[(count == 0) | (count < min_points)] = -1
    grid[dist <= clear_radius] = 0
    return grid.reshape(h, w)


@pytest.mark.parametrize("seed,box", [(1, (0.0, 15.0, -10.0, 10.0, -1.5, 10.0)), (2, (2.0, 9.0, -3.0, 7.5, 0.0, 4.0))])
def test_crop_against_numpy(seed, box):
    p = scan(60_000, seed)
    got, idx = O.crop(p, A.CropBox(*box))
    want, widx = numpy_crop(p, A.CropBox(*box))
    assert np.array_equal(idx, widx) and np.array_equal(got.view(np.uint32), want.view(np.uint32)) and 500 < idx.size < p.shape[0]
    assert not np.isin(np.arange(6), idx).any() or box[0] != 0.0      # points on the default box's faces are dropped


@pytest.mark.parametrize("seed,leaf,n", [(3, 0.2, 20_000), (4, 0.05, 8_000), (5, 1.0, 20_000)])
def test_voxel_grid_against_numpy(seed, leaf, n):
    p, _ = numpy_crop(scan(n, seed), A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0))
    got, keys, counts, _ = O.voxel_grid(p, leaf)
    want, wkeys = numpy_voxel_grid(p, leaf)
    assert got.shape == want.shape and got.shape[0] > 50 and counts.sum() == p.shape[0]
    assert np.array_equal(keys.astype(np.int64), wkeys)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))   # centroids bit for bit


def quat(yaw, pitch, roll=0.0):
    cy, sy, cp, sp, cr, sr = np.cos(yaw / 2), np.sin(yaw / 2), np.cos(pitch / 2), np.sin(pitch / 2), np.cos(roll / 2), np.sin(roll / 2)
    return np.array([sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy])


@pytest.mark.parametrize("seed,yaw,pitch,roll", [(6, 0.0, 0.4363, 0.0), (7, 2.6, 0.1245, 0.0), (8, -1.1, 0.3, 0.7)])
def test_transform_against_numpy(seed, yaw, pitch, roll):
    p = scan(30_000, seed)
    t, q = np.array([9.25, -3.0, 2.0]), quat(yaw, pitch, roll)
    got, want = O.transform(p, t, q), numpy_transform(p, t, q)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("seed,res,width,height,min_pts", [(9, 1.0, 100.0, 100.0, 10), (10, 0.25, 40.0, 30.0, 3), (11, 0.7, 33.0, 21.5, 1)])
def test_drivable_grid_against_numpy(seed, res, width, height, min_pts):
    rng = np.random.default_rng(seed)
    n = 120_000
    p = np.zeros((n, 3), f32)
    p[:, :2] = rng.uniform(-30, 30, (n, 2))
    p[:, 2] = 0.2 * np.sin(0.3 * p[:, 0]) - 2.0 + np.where(rng.random(n) < 0.15, rng.uniform(0, 2.5, n), 0.0)
    p[rng.choice(n, 30, replace=False), rng.integers(0, 3, 30)] = np.nan
    t, q = np.array([9.25, -3.0, 2.0]), quat(2.6, 0.1245)
    robot, start = (9.0, -3.0), (7.5, -1.0)
    got = O.drivable_area(p, t, q, robot, start, grid_resolution=res, map_width=width, map_height=height, min_points_per_cell=min_pts)
    want = numpy_drivable(p, t, q, robot, start, res, width, height, 0.3, min_pts, 3.0)
    assert got.shape == want.shape and np.array_equal(got, want)
    assert (got == 100).any() and (got == 0).any()
