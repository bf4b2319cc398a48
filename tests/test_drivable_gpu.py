"""calc_drivable_area's occupancy grid (plc_drivable_area) against the oracle's restatement of
calc_drivable_area.cpp:87-232, through the C ABI.  Integer/byte work: the bar is bit-exact grids."""
import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import _abi as A

pytestmark = pytest.mark.gpu


def quat_zyx(yaw, pitch):
    cy, sy, cp, sp = np.cos(yaw / 2), np.sin(yaw / 2), np.cos(pitch / 2), np.sin(pitch / 2)
    return np.array([-sy * sp, cy * sp, sy * cp, cy * cp])


def robot_scan(n, seed, spread=30.0):
    """Filtered robot scan in the sensor frame: rolling ground, a few walls and clumps, some non-finite points."""
    rng = np.random.default_rng(seed)
    xy = rng.uniform(-spread, spread, (n, 2))
    z = 0.2 * np.sin(0.3 * xy[:, 0]) * np.cos(0.2 * xy[:, 1]) + rng.normal(0, 0.02, n) - 2.0
    wall = rng.random(n) < 0.15
    z[wall] += rng.uniform(0, 2.5, int(wall.sum()))
    pts = np.column_stack([xy, z]).astype(np.float32)
    pts[rng.choice(n, 30, replace=False), rng.integers(0, 3, 30)] = np.nan
    pts[rng.choice(n, 5, replace=False), 0] = np.inf
    return pts


def params(**kw):
    p = A.DrivableParams()
    A.lib().plc_drivable_params_default(p)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


@pytest.mark.parametrize("n,res,width,height,min_pts,seed", [
    (200_000, 1.0, 100.0, 100.0, 10, 1),          # the node's defaults
    (300_000, 0.25, 40.0, 30.0, 3, 2),
    (50_000, 0.7, 33.0, 21.5, 1, 3),              # sizes that do not divide: (int) truncation of the grid size
    (1_000_000, 0.1, 60.0, 60.0, 5, 4),           # a config-4 sized scan on a fine grid
])
def test_grid_matches_oracle(ctx, n, res, width, height, min_pts, seed):
    pts = robot_scan(n, seed)
    t, q = np.array([9.25, -3.0, 2.0]), quat_zyx(2.6, 0.1245)
    robot, start = (9.0, -3.0), (7.5, -1.0)
    prm = params(grid_resolution=res, map_width=width, map_height=height, min_points_per_cell=min_pts)
    got, info = ctx.drivable_area(pts, t, q, robot, start, prm)
    ref = O.drivable_area(pts, t, q, robot, start, grid_resolution=res, map_width=width, map_height=height,
                          min_points_per_cell=min_pts)
    assert got.shape == ref.shape == (info.height, info.width)
    assert (info.origin_x, info.origin_y) == (robot[0] - width / 2.0, robot[1] - height / 2.0)
    assert np.array_equal(got, ref)
    assert set(np.unique(got)) <= {-1, 0, 100} and (got == 100).any() and (got == 0).any()
    assert min_pts == 1 or (got == -1).any()


def test_quirks_and_edges(ctx):
    # a point just below the map origin truncates to cell 0 (static_cast<int> rounds towards zero, :151-152);
    # cells inside start_clear_radius are free whatever they hold (:194-197); one point in a cell: gradient 0
    prm = params(grid_resolution=1.0, map_width=4.0, map_height=4.0, min_points_per_cell=1, start_clear_radius=0.6)
    robot = (2.0, 2.0)                                 # map origin (0, 0)
    pts = np.array([[-0.5, -0.5, 0.0], [-0.5, -0.5, 5.0],        # both land in cell (0, 0): gradient 5 -> 100
                    [-1.0, 0.5, 0.0],                            # (int)(-1.0) = -1: outside
                    [1.5, 1.5, 0.0], [1.5, 1.5, 0.3],            # gradient 0.3f > 0.3: float32 0.3 is above the double
                    [2.5, 1.5, 0.0], [2.5, 1.5, 0.25],           # gradient 0.25 <= 0.3: free
                    [3.5, 3.5, 1.0],                             # a single point: gradient 0 -> free
                    [3.99, 0.5, 0.0], [4.0, 0.5, 0.0],           # x = 4.0 is cell 4: outside
                    [0.5, 3.5, 9.0], [0.5, 3.5, -9.0]], np.float32)
    start = (0.5, 3.5)                                 # clears cell (0, 3) although its gradient is 18
    got, info = ctx.drivable_area(pts, None, None, robot, start, prm)
    ref = O.drivable_area(pts, None, None, robot, start, grid_resolution=1.0, map_width=4.0, map_height=4.0,
                          min_points_per_cell=1, start_clear_radius=0.6)
    assert np.array_equal(got, ref)
    assert got[0, 0] == 100 and got[1, 1] == 100 and got[1, 2] == 0 and got[3, 3] == 0 and got[3, 0] == 0
    assert got[0, 3] == 0 and got[0, 1] == -1 and got[2, 2] == -1
    # an empty scan publishes nothing (:108-112); a scan with no point inside the map publishes an unknown map
    none, info = ctx.drivable_area(np.zeros((0, 3), np.float32), None, None, robot, start, prm)
    assert none is None and info.published == 0 and (info.width, info.height) == (4, 4)
    assert O.drivable_area(np.zeros((0, 3), np.float32), None, None, robot, start, map_width=4.0, map_height=4.0) is None
    far = np.full((10, 3), 1e6, np.float32)
    got, _ = ctx.drivable_area(far, None, None, robot, (100.0, 100.0), prm)
    assert np.all(got == -1)
    # unaligned wire layout: point_step 19, x/y/z at 1/6/11
    rec = np.zeros((pts.shape[0], 19), np.uint8)
    raw = pts.view(np.uint8).reshape(-1, 12)
    rec[:, 1:5], rec[:, 6:10], rec[:, 11:15] = raw[:, 0:4], raw[:, 4:8], raw[:, 8:12]
    got2, _ = ctx.drivable_area(rec.reshape(-1), None, None, robot, start, prm, point_step=19, off=(1, 6, 11))
    assert np.array_equal(got2, ref)
    # capacity too small
    import ctypes as C
    v = A.cloud_view(pts)
    info = A.DrivableInfo()
    small = np.zeros(3, np.int8)
    st = A.lib().plc_drivable_area(ctx._h, C.byref(v), None, None, 2.0, 2.0, 0.0, 0.0, C.byref(prm), small.ctypes.data, 3,
                                   C.byref(info))
    assert st == A.PLC_E_CAPACITY and info.n_cells == 16
