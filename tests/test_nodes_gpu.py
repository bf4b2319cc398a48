"""Node-level drop-in tests: the ROS-free node cores (include/plc_node.h) driven the way the rclcpp
wrappers drive them — PointCloud2 buffers into the callbacks, a TF result, the timer callback — and
checked against the oracle's restatement of the same node (virtual_lidar.cpp:164-548)."""
import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import _abi as A, scenes
from pointcloud_processor_b200.nodes import VirtualLidarNode, filter_scan, merge_clouds
from test_sweep_gpu import assert_same_optimum

pytestmark = pytest.mark.gpu

F = A  # flag bit names live in the header; mirror them here
RANGE_F, FOV_F, VIS_F, RANGE_M, FOV_M, VIS_M = 1, 2, 4, 8, 16, 32


def xyzrgb_records(xyz):
    """PointXYZRGB wire layout (point_step 32, x/y/z at 0/4/8), as the upstream publishers send."""
    rec = np.zeros((xyz.shape[0], 8), np.float32)
    rec[:, :3] = xyz
    return rec


def colour_counts(flags, dual):
    rng = (flags & RANGE_F) > 0
    fov = (flags & FOV_F) > 0
    vis = (flags & VIS_F) > 0
    if dual:
        rng |= (flags & RANGE_M) > 0
        fov |= (flags & FOV_M) > 0
        vis |= (flags & VIS_M) > 0
    blue = ~rng
    yellow = rng & ~fov
    red = rng & fov & ~vis
    green = rng & fov & vis
    return dict(total=flags.size, green=int(green.sum()), red=int(red.sum()), blue=int(blue.sum()), yellow=int(yellow.sum()))


@pytest.mark.parametrize("shape,n_pts,num_candidates", [("L", 50_000, 100), ("rect", 30_000, 49)])
def test_virtual_lidar_node_cycle(ctx, shape, n_pts, num_candidates):
    terrain = scenes.terrain(shape, n_pts, seed=21, side=24.0)
    area, _ = scenes.excavation_area(shape, 0.05)
    node = VirtualLidarNode(ctx)
    node.set_parameter("num_candidates", num_candidates)
    # guard (:455): nothing happens until grid, terrain and TF are all there
    assert node.run_optimization().ran == 0
    node.excavation_area_callback(xyzrgb_records(area))
    node.terrain_callback(xyzrgb_records(terrain))
    assert node.run_optimization().ran == 0          # TF lookup still failing
    node.set_base_transform((0.0, 0.0, 0.0))
    res = node.run_optimization()
    assert res.ran == 1
    log = node.take_log()

    # grid: lattice positions identical to the oracle's generateExcavationGrid3D (defaults 0.1 m, 10 layers)
    cells, flags = node.cells()
    ocells, ob = O.generate_cells(area, O.compute_normals(area, 1.5), 0.1, 10, 1.5)
    assert cells.size == ocells.size == res.n_cells
    for k in ("x", "y", "z"):
        assert np.array_equal(cells[k], ocells[k])
    nrm = np.stack([cells["nx"], cells["ny"], cells["nz"]], 1)
    onrm = np.stack([ocells["nx"], ocells["ny"], ocells["nz"]], 1)
    assert np.abs(nrm - onrm).max() < 1e-3           # PCA normals: setup input, not parity-gated
    assert f"Generated 3D grid: {cells.size} valid cells across 10 vertical layers" in log

    # candidates: identical to the oracle's generateCandidatePositions for the same fixed pose
    fixed = np.zeros(1, A.POSE_DTYPE)
    fixed["x"], fixed["y"], fixed["z"], fixed["pitch"] = 0.4, 0.5, 3.5, -np.pi / 6
    cands = node.candidates()
    tix = O.Index(terrain, O.GRID)
    ocands = O.generate_candidates(tix, ob, fixed, num_candidates, 3.0, 1.1)
    assert cands.size == ocands.size == res.n_candidates
    for k in ("x", "y", "z"):
        assert np.array_equal(cands[k], ocands[k]), k
    for k in ("pitch", "yaw"):                       # device atan2 (2 ulp) against libm's
        assert np.allclose(cands[k], ocands[k], rtol=0, atol=4e-15), k

    # sweep: the oracle on the node's own cells and candidates
    ref = O.sweep(tix, None, cells, A.default_params(), fixed, cands)
    assert ref["best_index"] >= 0
    # the node reports a pose, not an index: find it, and require it to be the oracle's optimum
    # (or an exact tie with it: mirrored poses of the symmetric lattice, see assert_same_optimum)
    hit = np.flatnonzero((cands["x"] == res.best.x) & (cands["y"] == res.best.y) & (cands["z"] == res.best.z))
    assert hit.size == 1
    assert_same_optimum(int(hit[0]), ref)
    b = cands[int(hit[0])]
    assert res.best_score == pytest.approx(ref["best_score"], rel=1e-12)
    assert res.fixed_total == pytest.approx(ref["fixed_total"], rel=1e-12)
    assert np.array_equal(flags, ref["flags"])       # incl. the stale-flag carry-over of the candidate loop (Q3)
    for got, dual in ((res.zx120_only, False), (res.dual, True)):
        want = colour_counts(ref["flags"], dual)
        assert {k: getattr(got, k) for k in want} == want
    assert res.range_ok == int(((ref["flags"] & RANGE_F) > 0).sum())
    assert res.visibility_ok == int(((ref["flags"] & VIS_F) > 0).sum())
    # the operator-visible log block (:419-451, :522-543)
    assert "ZX120 LiDAR Only Evaluation" in log and "Dual LiDAR Configuration (ZX120 + Mobile)" in log
    assert f"Best Mobile LiDAR Position: ({b['x']:.2f}, {b['y']:.2f}, {b['z']:.2f})" in log
    assert f"Total Score: {ref['best_score']:.2f}" in log
    assert f"  Total cells: {cells.size}" in log

    # parameters are re-read at the start of every cycle (:457); an observed cloud switches the fixed sensor
    node.set_parameter("max_distance", 6.0)
    node.set_parameter("sensor_height", 1.5)
    obs = (area[::7] + np.float32(0.03)).astype(np.float32)
    node.zx120_points_callback(xyzrgb_records(obs))
    res2 = node.run_optimization()
    prm = A.default_params()
    prm.max_distance = 6.0
    cands2 = node.candidates()
    ocands2 = O.generate_candidates(tix, ob, fixed, num_candidates, 3.0, 1.5)
    assert np.array_equal(cands2["z"], ocands2["z"])
    flags_in = ref["flags"].copy()                   # the cells keep their flags between cycles
    ref2 = O.sweep(tix, O.Index(obs, O.GRID), cells, prm, fixed, cands2, flags_inout=flags_in)
    assert res2.best_score == pytest.approx(ref2["best_score"], rel=1e-12)
    assert np.array_equal(node.cells()[1], ref2["flags"])
    assert res2.observed_points == obs.shape[0]
    assert f"ZX120 point cloud updated: {obs.shape[0]} points" in node.take_log()
    node.close()


def test_node_callbacks_take_device_buffers(ctx):
    """Every plc_cloud_view::data may be device memory (plc.h): the node core must hand the pointer
    on without reading it on the host (StoredCloud used to memcpy from it)."""
    from pointcloud_processor_b200.context import DevBuffer
    terrain = xyzrgb_records(scenes.terrain("rect", 30_000, seed=22, side=24.0))
    area = xyzrgb_records(scenes.excavation_area("rect", 0.05)[0])
    obs = xyzrgb_records((area[::9, :3] + np.float32(0.02)).astype(np.float32))

    def cycle(make):
        node = VirtualLidarNode(ctx)
        node.set_parameter("num_candidates", 36)
        node.excavation_area_callback(make(area))
        node.terrain_callback(make(terrain))
        node.zx120_points_callback(make(obs))
        node.set_base_transform((0.0, 0.0, 0.0))
        res = node.run_optimization()
        out = (res.ran, res.best_score, res.fixed_total, (res.best.x, res.best.y, res.best.z), node.cells()[1].copy(),
               node.candidates().copy())
        node.close()
        return out

    keep = []

    def on_device(rec):
        b = DevBuffer(ctx, rec.nbytes).upload(rec)
        keep.append(b)
        return b.view(rec.shape[0], 32)

    host, dev = cycle(lambda rec: rec), cycle(on_device)
    assert host[0] == dev[0] == 1
    assert host[1] == dev[1] and host[2] == dev[2] and host[3] == dev[3]
    assert np.array_equal(host[4], dev[4]) and np.array_equal(host[5], dev[5])
    for b in keep:
        b.free()


def test_node_with_only_empty_terrain_messages(ctx):
    """Upstream's guard is `!terrain_cloud_` (:455): an EMPTY /excavated_terrain message makes the pointer non-null
    while no KD-tree exists, so the cycle runs with every gated ray visible (:721, :727) and ground height 0 (:601)."""
    area, _ = scenes.excavation_area("rect", 0.05)
    node = VirtualLidarNode(ctx)
    node.set_parameter("num_candidates", 49)
    node.excavation_area_callback(xyzrgb_records(area))
    node.set_base_transform((0.0, 0.0, 0.0))
    assert node.run_optimization().ran == 0                      # no terrain message at all: guard holds
    node.terrain_callback(np.zeros((0, 8), np.float32))
    res = node.run_optimization()
    assert res.ran == 1
    cells, flags = node.cells()
    cands = node.candidates()
    assert cands.size > 0 and np.all(cands["z"] == 0.0 + 1.1)    # getGroundHeight -> 0.0
    fixed = np.zeros(1, A.POSE_DTYPE)
    fixed["x"], fixed["y"], fixed["z"], fixed["pitch"] = 0.4, 0.5, 3.5, -np.pi / 6
    ref = O.sweep(None, None, cells, A.default_params(), fixed, cands)
    assert ref["counters"]["rays_visible"] == ref["counters"]["rays_gated_in"] > 0
    assert res.best_score == pytest.approx(ref["best_score"], rel=1e-12)
    assert np.array_equal(flags, ref["flags"])
    # a real terrain arrives, then another empty message: the index stays (upstream keeps its tree), heights go to 0
    terrain = scenes.terrain("rect", 30_000, seed=23, side=24.0)
    node.terrain_callback(xyzrgb_records(terrain))
    with_terrain = node.run_optimization()
    assert with_terrain.best_score < res.best_score              # occlusion now counts
    node.terrain_callback(np.zeros((0, 8), np.float32))
    node.run_optimization()
    assert np.all(node.candidates()["z"] == 1.1)
    node.close()


def test_filter_and_merger_nodes(ctx):
    from test_filter_merge_gpu import scan
    pts = scan(120_000, 8)
    out = filter_scan(ctx, pts)                      # defaults 15 / 10 / 10 / 0.2 (pointcloud_filter.cpp:30-39)
    box = A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0)
    cent = O.voxel_grid(O.crop(pts, box)[0], 0.2)[0]
    assert np.array_equal(out[:, :3].view(np.uint32), cent.view(np.uint32))
    assert filter_scan(ctx, np.zeros((0, 3), np.float32)).shape[0] == 0          # Q7: empty in, empty out

    assert out.shape[0] > 2000
    nr = out.shape[0] * 5 // 9
    nb = out.shape[0] - nr
    robot = np.ascontiguousarray(out[:nr])
    backhoe = np.ascontiguousarray(out[nr:])
    # static TFs of launch/tf_*.launch.xml:4 (pitch about y) composed with a map pose
    def quat_y(a):
        return np.array([0.0, np.sin(a / 2), 0.0, np.cos(a / 2)])
    tr, qr = np.array([3.25, -1.0, 2.0]), quat_y(0.1245)
    tb, qb = np.array([0.55, 0.4, 3.5]), quat_y(0.4363)
    merged, rc, bc = merge_clouds(ctx, dict(cloud=robot, t=tr, q=qr), dict(cloud=backhoe, t=tb, q=qb))
    assert merged.shape[0] == nr + nb and rc.shape[0] == nr and bc.shape[0] == nb
    xyz = lambda rec: rec[:, :12].copy().view(np.float32).reshape(-1, 3)
    rgb = lambda rec: rec[:, 16:20].copy().view(np.uint32).ravel() & 0xFFFFFF
    assert np.array_equal(xyz(merged)[:nr].view(np.uint32), O.transform(robot[:, :3], tr, qr).view(np.uint32))
    assert np.array_equal(xyz(merged)[nr:].view(np.uint32), O.transform(backhoe[:, :3], tb, qb).view(np.uint32))
    assert np.all(rgb(merged)[:nr] == 0xFF0000) and np.all(rgb(merged)[nr:] == 0x0000FF)   # robot red first, backhoe blue
    assert np.array_equal(rc, merged[:nr]) and np.array_equal(bc, merged[nr:])
    # a TF failure skips that source (:389-393); no message at all skips it too
    merged, rc, bc = merge_clouds(ctx, dict(cloud=robot, t=None, q=None), dict(cloud=backhoe, t=tb, q=qb))
    assert merged.shape[0] == nb and rc.shape[0] == 0 and np.all(rgb(merged) == 0x0000FF)
    merged, rc, bc = merge_clouds(ctx, None, None)
    assert merged.shape[0] == 0
