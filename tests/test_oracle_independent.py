"""An independent check of the oracle's sweep ("What's missing" 7): the visibility masks of
oracle/vl_sweep.cpp are recomputed here with NumPy + SciPy's cKDTree, which shares no code with oracle/
(three radius indices under one `ray_clear`) nor with the CUDA path.

What is restated, from /root/reference/src/virtual_lidar.cpp:
  :656-669  range gate  L in [MIN_DISTANCE, max_distance]
  :703-714  FOV gate    |atan2(dz, hypot(dx, dy)) - pitch| <= pi/2
  :754-800  sampled ray s_0 = 0.5, s += 0.3 while s < L - 0.08; query = float32(sensor + (d / L) * s);
            blocked iff a terrain point lies within 0.08 * 0.7 of the query

The tree answers in float64 on the float32 data, while FLANN's L2_Simple<float> rounds every operation to
float32, so a sample whose nearest point is within `BAND` of the search radius is not
decided by this check: rays that depend on such a sample are "grazing" and excluded (the count is asserted
to be a tiny fraction and printed).  Every other ray must agree with the oracle bit for bit.
"""
import numpy as np
import pytest

from oracle import binding as O

scipy_spatial = pytest.importorskip("scipy.spatial")

BAND = 1e-5          # metres either side of the search radius (the north star's grazing band)


def independent_masks(terrain_xyz, cells, prm, cands):
    """(visible, grazing, gated) boolean arrays of shape (C, M)."""
    pts = np.asarray(terrain_xyz[:, :3], np.float32)
    pts = pts[np.all(np.isfinite(pts), axis=1)].astype(np.float64)
    tree = scipy_spatial.cKDTree(pts)
    r = prm.vis_radius * prm.block_radius_scale
    cx, cy, cz = (cells[k].astype(np.float64) for k in ("x", "y", "z"))
    M = cells.size
    # the upstream accumulation of the sample parameter (:765, :796)
    s_tab = []
    s = prm.ray_start
    while s < prm.max_distance:
        s_tab.append(s)
        s += prm.ray_step
    s_tab = np.array(s_tab)
    vis = np.zeros((cands.size, M), bool)
    graz = np.zeros((cands.size, M), bool)
    gated = np.zeros((cands.size, M), bool)
    for i, p in enumerate(cands):
        d = np.stack([cx - p["x"], cy - p["y"], cz - p["z"]], 1)
        L = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])
        ok = (L >= prm.min_distance) & (L <= prm.max_distance)
        elev = np.arctan2(d[:, 2], np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]))
        ok &= np.abs(elev - p["pitch"]) <= prm.fov_vertical / 2.0
        gated[i] = ok
        with np.errstate(invalid="ignore", divide="ignore"):
            u = d / L[:, None]
        n = np.searchsorted(s_tab, L - prm.vis_radius, side="left")      # #{k : s_k < L - vis_radius}
        n = np.where(ok, n, 0)
        kmax = int(n.max()) if M else 0
        first_block = np.full(M, kmax + 1)     # index of the first certainly blocked sample
        first_graze = np.full(M, kmax + 1)     # index of the first undecided sample
        for k in range(kmax):
            live = np.nonzero(n > k)[0]
            q = (np.array([p["x"], p["y"], p["z"]]) + u[live] * s_tab[k]).astype(np.float32).astype(np.float64)
            dist, _ = tree.query(q, k=1, distance_upper_bound=r + 2 * BAND, workers=-1)
            blocked = dist < r - BAND
            undecided = (dist >= r - BAND) & (dist <= r + BAND)
            jb, jg = live[blocked], live[undecided]
            first_block[jb] = np.minimum(first_block[jb], k)
            first_graze[jg] = np.minimum(first_graze[jg], k)
        vis[i] = ok & (first_block > kmax)
        # an undecided sample matters only if it comes before the first certain block
        graz[i] = ok & (first_graze < first_block)
    return vis, graz, gated


def unpack(mask, m):
    bits = np.unpackbits(mask.view(np.uint8), axis=1, bitorder="little")
    return bits[:, :m].astype(bool)


@pytest.mark.parametrize("name,n_poses", [("tiny", None), ("cfg1", 12)])
def test_oracle_masks_match_a_scipy_restatement(scene_cache, name, n_poses):
    s = scene_cache(name)
    cands = s.candidates
    if n_poses is not None:
        sel = np.sort(np.random.default_rng(17).choice(cands.size, n_poses, replace=False))
        cands = cands[sel]
    ref = O.sweep(O.Index(s.terrain, O.GRID), None, s.cells, s.params, s.fixed, cands, threads=8)
    got_vis, graz, gated = independent_masks(s.terrain, s.cells, s.params, cands)
    ora_vis = unpack(ref["mask"], s.cells.size)
    pairs = got_vis.size
    n_graz = int(graz.sum())
    differ = (got_vis != ora_vis) & ~graz
    print(f"{name}: {pairs} rays, {int(gated.sum())} gated in, {n_graz} grazing (excluded), {int(differ.sum())} differ")
    assert int(gated.sum()) == ref["counters"]["rays_gated_in"]
    assert int(differ.sum()) == 0
    assert n_graz <= max(4, int(2e-3 * pairs))          # the band is 1e-5 m wide: a handful of rays at most
    assert 0 < int(got_vis.sum()) < int(gated.sum())   # the scene exercises both outcomes


def test_hand_scene_block_tunnel_and_radius():
    """One point on a sample, between two samples, and either side of the radius."""
    from pointcloud_processor_b200 import _abi as A
    prm = A.default_params()
    cells = np.zeros(1, A.CELL_DTYPE)
    cells["x"], cells["nz"] = 5.0, 1.0
    pose = np.zeros(1, A.POSE_DTYPE)
    pose["pitch"] = 0.0
    r = prm.vis_radius * prm.block_radius_scale
    for px, off, expect_visible in [(0.5 + 0.3 * 4, 0.0, False),       # on sample 4
                                    (0.5 + 0.3 * 4 + 0.15, 0.0, True),  # between samples: tunnels
                                    (0.5 + 0.3 * 4, r - 1e-4, False),
                                    (0.5 + 0.3 * 4, r + 1e-4, True)]:
        terrain = np.array([[px, off, 0.0], [100.0, 100.0, 100.0]], np.float32)
        vis, graz, gated = independent_masks(terrain, cells, prm, pose)
        ref = O.sweep(O.Index(terrain, O.BRUTE), None, cells, prm, pose[0:1], pose)
        assert gated[0, 0] and not graz[0, 0]
        assert vis[0, 0] == expect_visible
        assert bool(ref["mask"][0, 0] & 1) == expect_visible


# ---------------------------------------------------------------------------------------------------------------------
# The node's setup steps restated with cKDTree: the evaluation-cell lattice with its cell
# normals (/root/reference/src/virtual_lidar.cpp:236-340) and the candidate-pose lattice with its ground heights
# (:550-625).  PCL's radius search works on the float32-narrowed query in float32; the tree here answers in float64,
# so a decision that hangs on a point within BAND of a search radius is not checked (counted, asserted rare).
def independent_cells(area, normals, res, layers, normal_radius):
    pts = np.asarray(area[:, :3], np.float32).astype(np.float64)
    tree = scipy_spatial.cKDTree(pts)
    lo, hi = pts.min(axis=0) - res, pts.max(axis=0) + res                       # :239-254
    width = int(np.ceil((hi[0] - lo[0]) / res)) + 1
    height = int(np.ceil((hi[1] - lo[1]) / res)) + 1
    z_step = (hi[2] - lo[2]) / max(1, layers)
    i, j, k = np.meshgrid(np.arange(height), np.arange(width), np.arange(layers), indexing="ij")   # loop order :265-271
    x, y, z = lo[0] + j.ravel() * res, lo[1] + i.ravel() * res, lo[2] + k.ravel() * z_step + z_step / 2.0
    q = np.column_stack([x, y, z]).astype(np.float32).astype(np.float64)       # PointXYZRGB search point
    d, _ = tree.query(q, k=1)
    near, graze = d < res * 1.5, np.abs(d - res * 1.5) < BAND                   # :273, :289-299
    keep = near & ~graze
    out = dict(x=x, y=y, z=z, near=near, graze=graze, normals=np.zeros((x.size, 3)), normal_graze=np.zeros(x.size, bool))
    nrm = normals.astype(np.float64)
    finite = np.all(np.isfinite(nrm), axis=1)
    for c in np.flatnonzero(keep):                                              # :301-340
        idx = np.array(tree.query_ball_point(q[c], normal_radius + BAND), int)
        dd = np.linalg.norm(pts[idx] - q[c], axis=1)
        out["normal_graze"][c] = bool(np.any(np.abs(dd - normal_radius) < BAND))
        use = idx[(dd < normal_radius) & finite[idx]]
        s = nrm[use].sum(axis=0)
        n = np.linalg.norm(s)
        out["normals"][c] = s / n if use.size and n > 1e-6 else (0.0, 0.0, 1.0)  # GridCell's default normal (:31-35)
    return out, (lo, hi, width, height)


def test_cell_lattice_matches_independent_restatement():
    from pointcloud_processor_b200 import scenes
    area, analytic = scenes.excavation_area("L", 0.05)
    normals = O.compute_normals(area, 1.5)
    res, layers = 0.1, 10
    cells, b = O.generate_cells(area, normals, res, layers, 1.5)
    ind, (lo, hi, width, height) = independent_cells(area, normals, res, layers, 1.5)
    assert (b.width, b.height) == (width, height)
    assert (b.min_x, b.max_x, b.min_y, b.max_y, b.min_z, b.max_z) == (lo[0], hi[0], lo[1], hi[1], lo[2], hi[2])
    # membership: the oracle's cells are exactly the lattice points the tree calls near, outside the grazing band
    key = lambda x, y, z: set(zip(x.tolist(), y.tolist(), z.tolist()))
    mine_sure = key(*(ind[k][ind["near"] & ~ind["graze"]] for k in "xyz"))
    mine_maybe = key(*(ind[k][ind["graze"]] for k in "xyz"))
    theirs = key(cells["x"], cells["y"], cells["z"])
    assert mine_sure <= theirs <= (mine_sure | mine_maybe)
    assert len(mine_maybe) <= 1e-3 * len(theirs) and len(theirs) > 3000
    # order: the reference's i (y), j (x), k (z) loop order
    order = np.lexsort((cells["z"], cells["x"], cells["y"]))
    assert np.array_equal(order, np.arange(cells.size))
    # normals of the cells whose neighbour set does not hang on a grazing point
    pos = {p: c for c, p in enumerate(zip(ind["x"].tolist(), ind["y"].tolist(), ind["z"].tolist()))}
    at = np.array([pos[p] for p in zip(cells["x"].tolist(), cells["y"].tolist(), cells["z"].tolist())])
    ok = ~ind["normal_graze"][at] & ~ind["graze"][at]
    got = np.column_stack([cells["nx"], cells["ny"], cells["nz"]]).astype(np.float64)
    assert ok.mean() > 0.9          # lattice cells over a lattice-sampled area: exact-radius neighbours are common
    assert np.max(np.abs(got[ok] - ind["normals"][at][ok])) < 2e-7          # float32 storage of a double unit vector
    print(f"cells {cells.size}, grazing lattice points {len(mine_maybe)}, cells with a grazing neighbour {int((~ok).sum())}")


def independent_candidates(terrain, bounds, fixed_xy, num, search_radius, sensor_height):
    pts = np.asarray(terrain[:, :3], np.float32).astype(np.float64)
    tree2 = scipy_spatial.cKDTree(pts[:, :2])
    lo_x, hi_x, lo_y, hi_y = bounds.min_x, bounds.max_x, bounds.min_y, bounds.max_y
    ex = (lo_x - search_radius, hi_x + search_radius, lo_y - search_radius, hi_y + search_radius)
    cx, cy, cz = (lo_x + hi_x) / 2.0, (lo_y + hi_y) / 2.0, (bounds.min_z + bounds.max_z) / 2.0
    side = int(np.ceil(np.sqrt(float(num))))
    xs, ys = (ex[1] - ex[0]) / (side - 1), (ex[3] - ex[2]) / (side - 1)
    out, grazing = [], 0
    for i in range(side):
        for j in range(side):
            x, y = ex[0] + i * xs, ex[2] + j * ys
            if np.sqrt((x - fixed_xy[0]) ** 2 + (y - fixed_xy[1]) ** 2) < 0.5:                       # :570-572
                continue
            if lo_x <= x <= hi_x and lo_y <= y <= hi_y:                                              # :574
                continue
            # getGroundHeight (:599-625): points within 2 m of (x, y, 0) in 3-D AND within 1 m horizontally -> max z
            qf = np.array([np.float32(x), np.float32(y)], np.float64)
            idx = np.array(tree2.query_ball_point([x, y], 1.0 + 1e-3), int)
            g = 0.0
            if idx.size:
                dh = np.hypot(pts[idx, 0] - x, pts[idx, 1] - y)
                d3 = np.sqrt((pts[idx, 0] - qf[0]) ** 2 + (pts[idx, 1] - qf[1]) ** 2 + pts[idx, 2] ** 2)
                sure = (dh < 1.0 - BAND) & (d3 < 2.0 - BAND)
                maybe = (dh < 1.0 + BAND) & (d3 < 2.0 + BAND)
                zs, zm = pts[idx[sure], 2], pts[idx[maybe], 2]
                if zm.size and (not zs.size or zm.max() != zs.max()):
                    grazing += 1
                    out.append(None)
                    continue
                g = zs.max() if zs.size else 0.0
            z = g + sensor_height
            dx, dy, dz = cx - x, cy - y, cz - z
            hd = np.sqrt(dx * dx + dy * dy)
            if hd < 0.1:
                continue
            elev = np.arctan2(-dz, hd)
            if np.deg2rad(-85.0) <= elev <= np.deg2rad(85.0):
                out.append((x, y, z, -np.pi / 2 + elev, np.arctan2(dy, dx)))
    return out, grazing


def test_candidate_lattice_matches_independent_restatement():
    from pointcloud_processor_b200 import scenes
    area, _ = scenes.excavation_area("L", 0.05)
    terrain = scenes.terrain("L", 50_000, seed=21, side=24.0)
    _, b = O.generate_cells(area, O.compute_normals(area, 1.5), 0.1, 10, 1.5)
    fixed = np.zeros(1, O.A.POSE_DTYPE)
    fixed["x"], fixed["y"], fixed["z"] = 0.4, 0.5, 3.5
    got = O.generate_candidates(O.Index(terrain, O.GRID), b, fixed, 400, 3.0, 1.1)
    want, grazing = independent_candidates(terrain, b, (0.4, 0.5), 400, 3.0, 1.1)
    assert len(want) == got.size > 200 and grazing <= 2
    for g, w in zip(got, want):
        if w is None:
            continue
        assert (g["x"], g["y"], g["z"]) == w[:3]
        assert abs(g["pitch"] - w[3]) < 1e-12 and abs(g["yaw"] - w[4]) < 1e-12


# ---------------------------------------------------------------------------------------------------------------------
# The terrain generator's carve, restated from /root/reference/src/excavated_surface_generator.cpp:
#   :183-225  getTerrainHeight: 3-D radius search around (x, y, 0), mean z of the hits within the radius in 2-D, nearest
#             neighbour's z if there are none
#   :284-300  pit centre = zx120 transform * (offset_x, offset_y, 0), its z from getTerrainHeight, yaw from getRPY
#   :326-348  isInsideExcavationArea: the boxes grown by the slope offset at that depth
#   :457-476  processExcavation: a point stays unless it is inside
def independent_kept(pts32, t, q, prm):
    pts = pts32.astype(np.float64)
    ok = np.all(np.isfinite(pts), axis=1)
    tree = scipy_spatial.cKDTree(pts[ok])
    pz = pts[ok, 2]
    R = prm.terrain_search_radius

    def height(x, y):
        """(mean z, grazing) as getTerrainHeight forms it."""
        qf = np.array([np.float32(x), np.float32(y), 0.0], np.float64)
        idx = np.array(tree.query_ball_point(qf, R + BAND), int)
        if idx.size:
            p = tree.data[idx]
            d3 = np.linalg.norm(p - qf, axis=1)
            d2 = np.hypot(p[:, 0] - x, p[:, 1] - y)
            graze = bool(np.any(np.abs(d3 - R) < BAND) or np.any(np.abs(d2 - R) < BAND))
            use = (d3 < R) & (d2 <= R)
            if use.any():
                return float(pz[idx[use]].mean()), graze
            if graze:
                return 0.0, True
        _, j = tree.query(qf, k=1)
        return float(pz[j]), False

    x, y, z, w = q
    yaw = np.arctan2(2.0 * (w * z + x * y), 1.0 - 2.0 * (y * y + z * z))                 # getRPY's yaw of a unit quaternion
    rot = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w)], [2 * (x * y + z * w), 1 - 2 * (x * x + z * z)]])
    cxy = rot @ np.array([prm.offset_x, prm.offset_y]) + np.asarray(t[:2], float)       # z of the offset is 0: the tilt terms vanish for a yaw-only base
    cz, _ = height(cxy[0], cxy[1])
    if prm.l_shape_enabled:                                                             # getExcavationBoxes :137-181
        boxes = [(0.0, -prm.arm1_length / 2.0, prm.arm1_width, prm.arm1_length),
                 (prm.arm2_length / 2.0, -prm.arm1_length + prm.arm2_width / 2.0, prm.arm2_length, prm.arm2_width)]
    else:
        boxes = [(0.0, 0.0, prm.length, prm.width)]
    slope_offset = prm.depth / np.tan(np.deg2rad(prm.slope_angle_deg))
    grow_max = max(slope_offset, 0.0) + 1e-6

    def inside(xl, yl, zr):
        if zr < -prm.depth or zr > 0:
            return False
        cur = slope_offset * ((prm.depth + zr) / prm.depth)
        return any(abs(xl - bx) <= bl / 2.0 + cur and abs(yl - by) <= bw / 2.0 + cur for bx, by, bl, bw in boxes)

    kept, grazing = np.ones(pts.shape[0], bool), np.zeros(pts.shape[0], bool)
    c, s = np.cos(-yaw), np.sin(-yaw)
    for i in np.flatnonzero(ok):
        dx, dy = pts[i, 0] - cxy[0], pts[i, 1] - cxy[1]
        xl, yl = dx * c - dy * s, dx * s + dy * c
        if not any(abs(xl - bx) <= bl / 2.0 + grow_max and abs(yl - by) <= bw / 2.0 + grow_max for bx, by, bl, bw in boxes):
            continue                                                                    # outside the widest outline whatever the height
        h, g = height(pts[i, 0], pts[i, 1])
        zr = pts[i, 2] - h
        a, b_, c_ = inside(xl, yl, zr), inside(xl, yl, zr - 1e-6), inside(xl, yl, zr + 1e-6)
        kept[i] = not a
        grazing[i] = g or not (a == b_ == c_)
    return kept, grazing, (cxy[0], cxy[1], cz), yaw


@pytest.mark.parametrize("shape,yaw", [("L", 0.0), ("rect", 0.4)])
def test_generator_carve_matches_independent_restatement(shape, yaw):
    from pointcloud_processor_b200 import scenes, _abi as A
    rng = np.random.default_rng(8)
    n = 12_000
    pts = np.zeros((n, 3), np.float32)                                                   # rolling ground, un-carved, around the pit
    pts[:, :2] = rng.uniform(-2.0, 7.0, (n, 2))
    pts[:, 2] = 0.15 * np.sin(0.4 * pts[:, 0]) * np.cos(0.3 * pts[:, 1]) + rng.normal(0, 0.01, n) - rng.uniform(0, 1.2, n) * (rng.random(n) < 0.3)
    pts[rng.choice(n, 8, replace=False), rng.integers(0, 3, 8)] = np.nan
    prm = A.excavation_params(l_shape_enabled=1 if shape == "L" else 0)
    t = np.array([0.3, -0.2, 0.0])
    q = np.array([0.0, 0.0, np.sin(yaw / 2), np.cos(yaw / 2)])
    ns, na = A.excavation_counts(prm)
    terrain, area, info = O.excavate(pts, np.zeros(n, np.uint32), t, q, prm, ns, na)
    kept, grazing, centre, ind_yaw = independent_kept(pts, t, q, prm)
    assert abs(info.yaw - ind_yaw) < 1e-12
    assert abs(info.center[0] - centre[0]) < 1e-12 and abs(info.center[1] - centre[1]) < 1e-12 and abs(info.center[2] - centre[2]) < 1e-9
    got_xyz = np.ascontiguousarray(terrain[: info.n_kept, :12]).view(np.float32).reshape(-1, 3)
    # the oracle's kept points, in input order, are the independently kept ones, give or take the grazing ones
    sure_kept, sure_cut = kept & ~grazing, ~kept & ~grazing
    got_set = set(map(bytes, got_xyz.view(np.uint8).reshape(-1, 12)))
    as_bytes = lambda m: set(map(bytes, pts[m].view(np.uint8).reshape(-1, 12)))
    assert as_bytes(sure_kept) <= got_set and not (as_bytes(sure_cut) & got_set)
    assert grazing.sum() <= 0.01 * n and (~kept).sum() > 100 and info.n_kept < n
    order = [bytes(b) for b in pts[np.isin(np.arange(n), np.flatnonzero(kept | grazing))].view(np.uint8).reshape(-1, 12) if bytes(b) in got_set]
    assert order == [bytes(b) for b in got_xyz.view(np.uint8).reshape(-1, 12)]           # input order kept (:470-472)
    print(f"{shape}: {n} points, {int((~kept).sum())} carved, {int(grazing.sum())} grazing, kept {info.n_kept}")


# PCA normals of /excavation_area (/root/reference/src/virtual_lidar.cpp:209-234, pcl::NormalEstimation with a radius
# search): neighbours within the radius, covariance about their centroid, eigenvector of the smallest eigenvalue,
# flipped towards the viewpoint (0, 0, 0) by PCL and then to n_z >= 0 by the node.
def test_pca_normals_match_independent_restatement():
    from pointcloud_processor_b200 import scenes
    area, analytic = scenes.excavation_area("L", 0.05)
    got = O.compute_normals(area, 1.5).astype(np.float64)
    pts = np.asarray(area[:, :3], np.float32).astype(np.float64)
    tree = scipy_spatial.cKDTree(pts)
    rng = np.random.default_rng(4)
    checked = grazing = 0
    for i in rng.choice(pts.shape[0], 400, replace=False):
        idx = np.array(tree.query_ball_point(pts[i], 1.5 + BAND), int)
        d = np.linalg.norm(pts[idx] - pts[i], axis=1)
        if np.any(np.abs(d - 1.5) < BAND):
            grazing += 1
            continue
        nb = pts[idx[d < 1.5]]
        assert nb.shape[0] >= 3
        c = nb - nb.mean(axis=0)
        w, v = np.linalg.eigh(c.T @ c / nb.shape[0])
        n = v[:, 0]                                            # smallest eigenvalue first
        if np.dot(-pts[i], n) < 0:                             # flipNormalTowardsViewpoint, viewpoint (0, 0, 0)
            n = -n
        if n[2] < 0:                                           # :223-229 (the second flip decides the final sign)
            n = -n
        gap = w[1] - w[0]
        if gap < 1e-6 * max(w[2], 1e-12):                      # a degenerate neighbourhood has no defined normal
            continue
        assert abs(np.dot(n, got[i])) > 1 - 1e-6 and np.dot(n, got[i]) > 0, (i, n, got[i])
        checked += 1
    assert checked > 300
    print(f"normals: {checked} checked, {grazing} with a neighbour at the search radius")
