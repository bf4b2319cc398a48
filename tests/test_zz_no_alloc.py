"""plc_drivable_area_static — the entry point that allocates nothing (its per-cell scratch is a static __device__
array inside libplc.so, no counterpart in the reference) — against
plc_drivable_area, through the C ABI.  Byte work: the two grids must be identical, on two calls in a row that share
the scratch array.  Collected last on purpose (pytest -x stops at the first failure)."""
import numpy as np
import pytest

from pointcloud_processor_b200 import _abi as A
from pointcloud_processor_b200.context import DevBuffer

pytestmark = pytest.mark.gpu

MAX_CELLS = 1 << 20          # PLC_DRIVABLE_STATIC_MAX_CELLS (include/plc.h)


def quat_zyx(yaw, pitch):
    cy, sy, cp, sp = np.cos(yaw / 2), np.sin(yaw / 2), np.cos(pitch / 2), np.sin(pitch / 2)
    return np.array([-sy * sp, cy * sp, sy * cp, cy * cp])


def robot_scan(n, seed, spread):
    rng = np.random.default_rng(seed)
    xy = rng.uniform(-spread, spread, (n, 2))
    z = 0.2 * np.sin(0.3 * xy[:, 0]) * np.cos(0.2 * xy[:, 1]) + rng.normal(0, 0.02, n) - 2.0
    wall = rng.random(n) < 0.15
    z[wall] += rng.uniform(0, 2.5, int(wall.sum()))
    pts = np.column_stack([xy, z]).astype(np.float32)
    pts[rng.choice(n, 20, replace=False), rng.integers(0, 3, 20)] = np.nan
    return pts


def params(**kw):
    p = A.DrivableParams()
    A.lib().plc_drivable_params_default(p)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def static_grid(ctx, pts, t, q, robot, start, prm, scan_buf, grid_buf):
    """One call of the allocation-free entry point on caller-owned device buffers."""
    scan_buf.upload(pts)
    grid_buf.upload(np.full(grid_buf.nbytes, 77, np.int8))                # stale bytes the call must overwrite
    info = ctx.drivable_area_static(scan_buf.view(len(pts), 12), t, q, robot, start, prm, grid_buf.ptr, grid_buf.nbytes)
    ctx.sync()
    assert info.published == 1
    return grid_buf.download(info.n_cells, np.int8).reshape(info.height, info.width), info


def test_static_scratch_variant_equals_allocating_entry_point(ctx):
    t, q = np.array([9.25, -3.0, 2.0]), quat_zyx(2.6, 0.1245)
    robot, start = (9.0, -3.0), (7.5, -1.0)
    scan_buf, grid_buf = DevBuffer(ctx, 600_000 * 12), DevBuffer(ctx, MAX_CELLS)
    # two calls in a row on the shared scratch: a large fine grid first, then a smaller, coarser one whose cells
    # reuse the first call's leftovers (it would inherit counts and extrema if the reset were missing)
    cases = [(600_000, 0.1, 60.0, 60.0, 5, 11, 30.0),
             (150_000, 1.0, 100.0, 100.0, 10, 12, 50.0),
             (300_000, 0.0625, 64.0, 64.0, 2, 13, 32.0)]                  # 1024 x 1024 = exactly the maximum
    for n, res, width, height, min_pts, seed, spread in cases:
        pts = robot_scan(n, seed, spread)
        prm = params(grid_resolution=res, map_width=width, map_height=height, min_points_per_cell=min_pts)
        want, winfo = ctx.drivable_area(pts, t, q, robot, start, prm)
        got, info = static_grid(ctx, pts, t, q, robot, start, prm, scan_buf, grid_buf)
        assert (info.width, info.height, info.origin_x, info.origin_y) == (winfo.width, winfo.height, winfo.origin_x,
                                                                           winfo.origin_y)
        assert got.shape == want.shape
        assert np.array_equal(got, want)
        assert (got == 100).any() and (got == 0).any() and (got == -1).any()
    # the same input twice: identical grids, and no map-frame transform (t = q = NULL)
    pts = robot_scan(100_000, 14, 40.0)
    prm = params(grid_resolution=0.5, map_width=80.0, map_height=70.0, min_points_per_cell=3)
    want, _ = ctx.drivable_area(pts, None, None, (0.0, 0.0), (1.0, 1.0), prm)
    a, _ = static_grid(ctx, pts, None, None, (0.0, 0.0), (1.0, 1.0), prm, scan_buf, grid_buf)
    b, _ = static_grid(ctx, pts, None, None, (0.0, 0.0), (1.0, 1.0), prm, scan_buf, grid_buf)
    assert np.array_equal(a, want) and np.array_equal(b, want)
    scan_buf.free()
    grid_buf.free()


def test_static_scratch_limits_and_pointer_kinds(ctx):
    pts = robot_scan(1000, 15, 10.0)
    scan_buf, grid_buf = DevBuffer(ctx, pts.nbytes).upload(pts), DevBuffer(ctx, 4 * MAX_CELLS)
    view = scan_buf.view(len(pts), 12)
    # one row more than the maximum: refused with a message that names the limit, nothing written
    grid_buf.upload(np.full(4096, 55, np.int8))
    prm = params(grid_resolution=0.0625, map_width=64.0, map_height=64.0625)      # 1024 x 1025 cells
    with pytest.raises(A.PlcError) as e:
        ctx.drivable_area_static(view, None, None, (0.0, 0.0), (0.0, 0.0), prm, grid_buf.ptr, grid_buf.nbytes)
    assert e.value.status == A.PLC_E_CAPACITY and "1048576" in str(e.value) and "1049600" in str(e.value)
    ctx.sync()
    assert (grid_buf.download(4096, np.int8) == 55).all()
    # a caller's capacity below the grid: PLC_E_CAPACITY as in plc_drivable_area
    prm = params(grid_resolution=1.0, map_width=10.0, map_height=10.0)
    with pytest.raises(A.PlcError) as e:
        ctx.drivable_area_static(view, None, None, (0.0, 0.0), (0.0, 0.0), prm, grid_buf.ptr, 99)
    assert e.value.status == A.PLC_E_CAPACITY
    # host memory cannot be staged without allocating: refused, for the input and for the output
    host_out = np.zeros(100, np.int8)
    with pytest.raises(A.PlcError) as e:
        ctx.drivable_area_static(view, None, None, (0.0, 0.0), (0.0, 0.0), prm, host_out.ctypes.data, 100)
    assert e.value.status == A.PLC_E_INVALID_ARG
    with pytest.raises(A.PlcError) as e:
        ctx.drivable_area_static(A.cloud_view(pts), None, None, (0.0, 0.0), (0.0, 0.0), prm, grid_buf.ptr, 100)
    assert e.value.status == A.PLC_E_INVALID_ARG
    # an empty scan publishes nothing, as in plc_drivable_area
    empty = scan_buf.view(0, 12)
    info = ctx.drivable_area_static(empty, None, None, (0.0, 0.0), (0.0, 0.0), prm, grid_buf.ptr, 100)
    assert info.published == 0
    # and the context is still good for a normal call
    got, info = ctx.drivable_area(pts, None, None, (0.0, 0.0), (0.0, 0.0), prm)
    assert got.shape == (10, 10)
    scan_buf.free()
    grid_buf.free()
