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
