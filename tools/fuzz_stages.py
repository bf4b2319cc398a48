"""Randomised differential test of the node stages either side of the sweep: the CUDA path
through the C ABI against the oracle on random clouds, parameters, transforms and wire layouts.

    python tools/fuzz_stages.py [iterations] [seed]          (on a GPU box)

Prints one summary line; exits non-zero on the first deviation beyond the stated bars (bit-identical sets,
positions, order, colours and grids; float32 normal components and generated z within one ulp)."""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, os.fspath(ROOT))
sys.path.insert(0, os.fspath(ROOT / "tests"))
from oracle import binding as O  # noqa: E402
from pointcloud_processor_b200 import Context, _abi as A, scenes  # noqa: E402


def quat(rng, yaw_only=False):
    if yaw_only:
        a = rng.uniform(-np.pi, np.pi)
        return np.array([0.0, 0.0, np.sin(a / 2), np.cos(a / 2)])
    q = rng.normal(size=4)
    return q / np.linalg.norm(q)


def records(xyz, step, offs, rng):
    """xyz packed into PointCloud2-like records of `step` bytes with x/y/z at `offs` (other bytes random)."""
    n = xyz.shape[0]
    rec = rng.integers(0, 256, (n, step), dtype=np.uint8)
    raw = np.ascontiguousarray(xyz, np.float32).view(np.uint8).reshape(n, 12)
    for k, o in enumerate(offs):
        rec[:, o:o + 4] = raw[:, 4 * k:4 * k + 4]
    return rec.reshape(-1)


def layout(rng):
    step = int(rng.choice([12, 16, 19, 22, 32, 48]))
    if step == 12:
        return 12, (0, 4, 8)
    while True:        # a greedy pick of byte offsets can leave no room for the third field in a short record: draw again
        slots = rng.permutation(np.arange(0, step - 3, 4 if step % 4 == 0 and rng.random() < 0.7 else 1))
        chosen = []
        for s in slots:
            if all(abs(int(s) - c) >= 4 for c in chosen):
                chosen.append(int(s))
            if len(chosen) == 3:
                return step, tuple(chosen)


def one_drivable(ctx, rng):
    n = int(rng.integers(1, 40_000))
    pts = np.column_stack([rng.uniform(-25, 25, (n, 2)), rng.normal(-1.5, 0.4, n)]).astype(np.float32)
    if n > 50:
        pts[rng.choice(n, 5, replace=False), rng.integers(0, 3, 5)] = rng.choice([np.nan, np.inf, -np.inf], 5)
    t, q = rng.uniform(-5, 5, 3), quat(rng, yaw_only=rng.random() < 0.5)
    prm = A.DrivableParams()
    A.lib().plc_drivable_params_default(prm)
    prm.grid_resolution = float(rng.choice([0.25, 0.5, 1.0, 0.7, 1.3]))
    prm.map_width, prm.map_height = float(rng.uniform(5, 60)), float(rng.uniform(5, 60))
    prm.max_gradient = float(rng.uniform(0.05, 1.0))
    prm.min_points_per_cell = int(rng.integers(0, 12))
    prm.start_clear_radius = float(rng.uniform(0, 6))
    robot, start = rng.uniform(-3, 3, 2), rng.uniform(-6, 6, 2)
    use_tf = rng.random() < 0.8
    step, offs = layout(rng)
    got, info = ctx.drivable_area(records(pts, step, offs, rng), t if use_tf else None, q if use_tf else None, robot, start, prm,
                                  point_step=step, off=offs)
    ref = O.drivable_area(pts, t if use_tf else None, q if use_tf else None, robot, start, grid_resolution=prm.grid_resolution,
                          map_width=prm.map_width, map_height=prm.map_height, max_gradient=prm.max_gradient,
                          min_points_per_cell=prm.min_points_per_cell, start_clear_radius=prm.start_clear_radius)
    assert np.array_equal(got, ref), "drivable grid differs"
    return n


def one_generator_and_setup(ctx, rng):
    from test_excavate_gpu import assert_clouds_match
    l_shape = int(rng.random() < 0.6)
    prm = A.excavation_params(l_shape_enabled=l_shape, point_density=float(rng.choice([0.05, 0.08, 0.1, 0.15])),
                              depth=float(rng.uniform(0.4, 1.5)), slope_angle_deg=float(rng.uniform(50, 85)),
                              terrain_search_radius=float(rng.uniform(0.2, 0.8)), offset_x=float(rng.uniform(2, 5)),
                              offset_y=float(rng.uniform(-2, 2)), arm1_length=float(rng.uniform(1.2, 2.5)),
                              arm1_width=float(rng.uniform(0.8, 1.4)), arm2_length=float(rng.uniform(1.2, 2.5)),
                              arm2_width=float(rng.uniform(0.8, 1.4)), length=float(rng.uniform(1.0, 2.5)),
                              width=float(rng.uniform(0.8, 1.6)))
    ns, na = A.excavation_counts(prm)
    n = int(rng.integers(0, 25_000)) if rng.random() < 0.9 else 0
    z0 = float(rng.choice([0.0, 0.0, 0.2, -0.3, 2.5]))          # 2.5: every height through the nearest-neighbour fallback
    base_t, base_q = rng.uniform(-2, 2, 3), quat(rng, yaw_only=True)
    yaw = 2 * np.arctan2(base_q[2], base_q[3])
    c = base_t[:2] + np.array([np.cos(yaw) * prm.offset_x - np.sin(yaw) * prm.offset_y, np.sin(yaw) * prm.offset_x + np.cos(yaw) * prm.offset_y])
    xy = rng.uniform(c - 7, c + 7, (n, 2))
    z = z0 + 0.15 * np.sin(0.4 * xy[:, 0]) * np.cos(0.3 * xy[:, 1]) + rng.uniform(-0.02, 0.02, n)
    rec = np.zeros((n, 8), np.float32)
    rec[:, :2], rec[:, 2], rec[:, 3] = xy, z, 1.0
    colours = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    rec.view(np.uint32)[:, 4] = colours
    if n > 100:
        rec[rng.choice(n, 6, replace=False), rng.integers(0, 3, 6)] = np.nan
    ref_t, ref_a, ri = O.excavate(rec[:, :3], colours, base_t, base_q, prm, ns, na)
    got_t, got_a, gi = ctx.excavate(rec, base_t, base_q, prm)
    assert (gi.n_kept, gi.n_surface, gi.n_area) == (ri.n_kept, ri.n_surface, ri.n_area), "generator counts differ"
    dz = assert_clouds_match(got_t, ref_t, int(ri.n_kept), "terrain") + assert_clouds_match(got_a, ref_a, 0, "area")
    # setup on the generator's outputs
    area_xyz = np.ascontiguousarray(ref_a[:, :12]).view(np.float32).reshape(-1, 3)
    terrain_xyz = np.ascontiguousarray(ref_t[:, :12]).view(np.float32).reshape(-1, 3)
    res, layers = float(rng.choice([0.1, 0.15, 0.2, 0.3])), int(rng.integers(1, 9))
    normals = O.compute_normals(area_xyz, 1.5) if rng.random() < 0.5 else None
    ref_c, rb = O.generate_cells(area_xyz, normals, res, layers, 1.5)
    got_c, gb = ctx.generate_cells(ref_a.reshape(-1), normals, res, layers, 1.5, point_step=32)
    assert got_c.size == ref_c.size, "cell count differs"
    for k in ("x", "y", "z"):
        assert np.array_equal(got_c[k], ref_c[k]), "cell positions differ"
    dn = 0
    for k in ("nx", "ny", "nz"):
        d = np.abs(got_c[k].astype(np.float64) - ref_c[k])
        assert d.max(initial=0) <= 1.2e-7, "cell normals differ"
        dn += int(np.count_nonzero(d))
    for f in ("min_x", "max_x", "min_y", "max_y", "min_z", "max_z", "width", "height"):
        assert getattr(gb, f) == getattr(rb, f), "bounds differ"
    num, radius, height = int(rng.integers(2, 500)), float(rng.uniform(1, 6)), float(rng.uniform(0.5, 3))
    fixed = scenes.fixed_pose()
    ref_p = O.generate_candidates(O.Index(terrain_xyz, O.GRID), rb, fixed, num, radius, height)
    got_p = ctx.generate_candidates(ref_t.reshape(-1), rb, fixed, num, radius, height, point_step=32)
    assert got_p.size == ref_p.size, "candidate count differs"
    for k in ("x", "y", "z"):
        assert np.array_equal(got_p[k], ref_p[k]), "candidate positions differ"
    for k in ("pitch", "yaw"):
        assert np.allclose(got_p[k], ref_p[k], rtol=0, atol=4e-15), "candidate angles differ"
    return n, ref_c.size, ref_p.size, dz, dn


def run(ctx, iterations, seed):
    rng = np.random.default_rng(seed)
    tot = dict(scans=0, scan_points=0, clouds=0, cloud_points=0, cells=0, poses=0, z_last_bit=0, normal_last_bit=0)
    for _ in range(iterations):
        tot["scan_points"] += one_drivable(ctx, rng)
        tot["scans"] += 1
        n, c, p, dz, dn = one_generator_and_setup(ctx, rng)
        tot["clouds"] += 1
        tot["cloud_points"] += n
        tot["cells"] += c
        tot["poses"] += p
        tot["z_last_bit"] += dz
        tot["normal_last_bit"] += dn
    return tot


if __name__ == "__main__":
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    with Context(0) as ctx:
        print("fuzz_stages ok:", run(ctx, iters, seed))
