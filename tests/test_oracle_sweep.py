"""CPU tests of the sweep restatement as a whole: index kinds agree, the timing-faithful loop gives
the same answers, first-maximum argmax, stale-flag carry-over, fixed-sensor branches."""
import numpy as np

from oracle import binding as O
from pointcloud_processor_b200 import _abi as A


def test_sweep_same_for_every_index_and_mode(scene_cache):
    s = scene_cache("tiny")
    base = O.sweep(O.Index(s.terrain, O.BRUTE), None, s.cells, s.params, s.fixed, s.candidates)
    assert base["counters"]["rays"] == s.cells.size * s.candidates.size
    assert 0 < base["counters"]["rays_visible"] < base["counters"]["rays"]
    for kind in (O.GRID, O.KDTREE):
        for kw in (dict(), dict(faithful=True), dict(threads=4)):
            r = O.sweep(O.Index(s.terrain, kind), None, s.cells, s.params, s.fixed, s.candidates, **kw)
            assert np.array_equal(r["mask"], base["mask"])
            assert np.array_equal(r["scores"], base["scores"])
            assert np.array_equal(r["covered"], base["covered"])
            assert r["best_index"] == base["best_index"] and r["fixed_total"] == base["fixed_total"]
            assert r["counters"]["samples"] == base["counters"]["samples"]
            if kw.get("threads", 1) == 1:
                assert np.array_equal(r["flags"], base["flags"])


def test_argmax_keeps_first_maximum(scene_cache):
    s = scene_cache("tiny")
    ix = O.Index(s.terrain, O.GRID)
    r = O.sweep(ix, None, s.cells, s.params, s.fixed, s.candidates)
    b = r["best_index"]
    # duplicate the best candidate in front and behind: the first copy must win (strict '>', :471)
    dup = np.concatenate([s.candidates[b:b + 1], s.candidates, s.candidates[b:b + 1]])
    r2 = O.sweep(ix, None, s.cells, s.params, s.fixed, dup)
    assert r2["best_index"] == 0 and r2["best_score"] == r["best_score"]
    # no candidates: best_index -1, -inf (the node then publishes the default pose, Q5)
    r3 = O.sweep(ix, None, s.cells, s.params, s.fixed, s.candidates[:0])
    assert r3["best_index"] == -1 and r3["best_score"] == -np.inf


def test_stale_flags_follow_the_candidate_loop(scene_cache):
    s = scene_cache("tiny")
    ix = O.Index(s.terrain, O.GRID)
    # make the last candidate out of range for every cell: its range flag is written (false) but
    # the FOV/visibility flags keep the values of earlier candidates
    far = s.candidates[-1:].copy()
    far["x"] += 500.0
    cands = np.concatenate([s.candidates, far])
    r = O.sweep(ix, None, s.cells, s.params, s.fixed, cands)
    f = r["flags"]
    assert not np.any(f & A.FLAG_RANGE_MOBILE)
    prev = O.sweep(ix, None, s.cells, s.params, s.fixed, s.candidates)["flags"]
    keep = A.FLAG_FOV_MOBILE | A.FLAG_VIS_MOBILE | A.FLAG_RANGE_FIXED | A.FLAG_FOV_FIXED | A.FLAG_VIS_FIXED
    assert np.array_equal(f & keep, prev & keep)
    # flags persist across sweeps: a second sweep with only the far candidate leaves them in place
    flags = prev.copy()
    O.sweep(ix, None, s.cells, s.params, s.fixed, far, flags_inout=flags)
    assert np.array_equal(flags & keep, prev & keep) and not np.any(flags & A.FLAG_RANGE_MOBILE)


def test_fixed_sensor_branches(scene_cache):
    s = scene_cache("tiny")
    ix = O.Index(s.terrain, O.GRID)
    ray = O.sweep(ix, None, s.cells, s.params, s.fixed, s.candidates[:1])
    # observed cloud present: a cell with an observed point within 0.24 m is visible without a ray
    obs_pts = np.stack([s.cells["x"], s.cells["y"], s.cells["z"]], 1).astype(np.float32)[::3]
    obs = O.Index(obs_pts, O.GRID)
    with_obs = O.sweep(ix, obs, s.cells, s.params, s.fixed, s.candidates[:1])
    vis_ray = (ray["flags"] & A.FLAG_VIS_FIXED) != 0
    vis_obs = (with_obs["flags"] & A.FLAG_VIS_FIXED) != 0
    assert np.all(vis_obs[::3]) and np.all(vis_obs >= vis_ray) and vis_obs.sum() > vis_ray.sum()
    assert with_obs["counters"]["fixed_samples"] < ray["counters"]["fixed_samples"]
    # memo variant: observed cloud only, no cloud => nothing visible from the fixed sensor
    memo = A.memo_params()
    none = O.sweep(ix, None, s.cells, memo, s.fixed, s.candidates[:1])
    assert none["fixed_total"] == 0.0 and not np.any(none["flags"] & A.FLAG_VIS_FIXED)
    # no terrain at all: every gated ray is visible (:721, :727)
    open_air = O.sweep(None, None, s.cells, s.params, s.fixed, s.candidates)
    assert open_air["counters"]["rays_visible"] == open_air["counters"]["rays_gated_in"]
    assert open_air["counters"]["samples"] == 0


def test_terrain_point_order_is_irrelevant(scene_cache):
    s = scene_cache("tiny")
    a = O.sweep(O.Index(s.terrain, O.GRID), None, s.cells, s.params, s.fixed, s.candidates)
    perm = np.random.default_rng(1).permutation(s.terrain.shape[0])
    b = O.sweep(O.Index(s.terrain[perm], O.KDTREE), None, s.cells, s.params, s.fixed, s.candidates)
    assert np.array_equal(a["mask"], b["mask"]) and np.array_equal(a["scores"], b["scores"])
