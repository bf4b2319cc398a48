"""CPU tests of the product's boundary and host-side logic (no compute calls that need a GPU):
the C-ABI library loads and exports every symbol the headers declare, parameter presets match the
reference constants, and the host setup steps (cell lattice, candidates, normals) agree with the
oracle's restatement of virtual_lidar.cpp:209-340, 550-625."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from oracle import binding as O
from pointcloud_processor_b200 import _abi as A
from pointcloud_processor_b200 import context as ctxmod
from pointcloud_processor_b200 import scenes

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    names = set()
    for h in ("plc.h", "plc_host.h", "plc_node.h"):
        text = (ROOT / "include" / h).read_text()
        names |= set(re.findall(r"\b(plc_[a-z0-9_]+)\s*\(", text))
    return names - {"plc_status"}   # the enum is mentioned as "plc_status (0 = OK ...)" in a comment


def test_library_exports_every_declared_symbol():
    L = A.lib()
    declared = declared_symbols()
    assert declared == set(A.EXPORTS), declared ^ set(A.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert L.plc_abi_version() == 1
    assert A.strerror(A.PLC_OK) == "ok" and "terrain" in A.strerror(A.PLC_E_NO_TERRAIN)


def test_pod_layouts_match_the_header():
    assert C.sizeof(A.CloudView) == 32 and C.sizeof(A.Params) == 96
    assert C.sizeof(A.Counters) == 48 and C.sizeof(A.GridBounds) == 56
    assert A.CELL_DTYPE.itemsize == 40 and A.POSE_DTYPE.itemsize == 40
    assert C.sizeof(A.SweepOut) == 8 + 8 + 4 + 4 + 8 + 48 + 6 * 8


def test_parameter_presets_match_reference_constants():
    p = A.default_params()    # virtual_lidar.cpp:66-71, 100-114
    assert (p.max_distance, p.min_distance, p.ray_step, p.ray_start) == (15.0, 0.5, 0.3, 0.5)
    assert (p.vis_radius, p.block_radius_scale, p.observed_radius_scale) == (0.08, 0.7, 3.0)
    assert p.fov_vertical == 180.0 * np.pi / 180.0 and p.fov_horizontal < 0 and p.fixed_mode == 0
    m = A.memo_params()       # virtual_lidar_memo.cpp:102-113, 684-697
    assert (m.min_distance, m.ray_step, m.vis_radius, m.fixed_mode) == (1.0, 0.2, 0.05, 1)
    assert m.fov_vertical == 90.0 * np.pi / 180.0 and m.fov_horizontal == 180.0 * np.pi / 180.0


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(A.PlcError) as e:
        ctxmod.Context(0)
    assert e.value.status == A.PLC_E_CUDA


@pytest.mark.parametrize("shape,res,layers", [("L", 0.1, 10), ("rect", 0.1, 10), ("L", 0.2, 3)])
def test_cell_lattice_matches_oracle(shape, res, layers):
    area, normals = scenes.excavation_area(shape, 0.05)
    got, gb = ctxmod.host_generate_cells(area, normals, res, layers, 1.5)
    ref, rb = O.generate_cells(area, normals, res, layers, 1.5)
    assert got.size == ref.size and got.size > 0
    assert np.array_equal(got.view(np.uint8), ref.view(np.uint8))      # bit-exact incl. normals
    for f in ("min_x", "max_x", "min_y", "max_y", "min_z", "max_z", "width", "height"):
        assert getattr(gb, f) == getattr(rb, f)
    n = np.stack([got["nx"], got["ny"], got["nz"]], 1).astype(np.float64)
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-6) and np.all(got["nz"] >= 0)


def test_reference_default_sizes():
    # L-shape area cloud 5 575 points, lattice 35 x 29 x 10 at the node's defaults
    area, _ = scenes.excavation_area("L", 0.05)
    assert area.shape[0] == 5575
    cells, b = ctxmod.host_generate_cells(area, None, 0.1, 10, 1.5)
    assert (b.width, b.height) == (35, 29)
    assert 3000 < cells.size < 4500
    area_r, _ = scenes.excavation_area("rect", 0.05)
    assert area_r.shape[0] == 3130


def test_candidates_match_oracle(scene_cache):
    s = scene_cache("small")
    ix = O.Index(s.terrain, O.GRID)
    for num, height in ((36, 1.1), (100, 1.5), (7, 1.1)):
        got = ctxmod.host_generate_candidates(s.terrain, s.bounds, s.fixed, num, 3.0, height)
        ref = O.generate_candidates(ix, s.bounds, s.fixed, num, 3.0, height)
        assert got.size == ref.size > 0
        assert np.array_equal(got.view(np.uint8), ref.view(np.uint8))
    # lattice points inside the (margin-expanded) grid bbox are excluded (:575)
    inside = (got["x"] >= s.bounds.min_x) & (got["x"] <= s.bounds.max_x) & \
             (got["y"] >= s.bounds.min_y) & (got["y"] <= s.bounds.max_y)
    assert not inside.any()


def test_candidate_lattice_of_one_is_rejected():
    area, _ = scenes.excavation_area("rect", 0.05)
    _, b = ctxmod.host_generate_cells(area, None, 0.1, 10, 1.5)
    with pytest.raises(A.PlcError):      # g = 1 divides by zero upstream (:563)
        ctxmod.host_generate_candidates(area, b, scenes.fixed_pose(), 1, 3.0, 1.1)


def test_pca_normals_close_to_oracle_and_analytic():
    area, analytic = scenes.excavation_area("rect", 0.05)
    got = ctxmod.host_compute_normals(area, 0.3)
    ref = O.compute_normals(area, 0.3)
    ok = np.isfinite(got).all(1) & np.isfinite(ref).all(1)
    assert ok.mean() > 0.99
    # same plane fit up to eigen-solver round-off (sign fixed by the n_z >= 0 rule)
    dots = np.abs(np.sum(got[ok].astype(np.float64) * ref[ok], 1))
    assert np.percentile(dots, 1) > 0.999
    # bottom points away from the walls: the fitted normal is the analytic (0,0,1)
    bottom = (analytic[:, 2] == 1.0)
    assert np.median(got[bottom, 2]) > 0.99


def test_empty_area_gives_no_cells():
    cells, _ = ctxmod.host_generate_cells(np.zeros((0, 3), np.float32), None, 0.1, 10, 1.5)
    assert cells.size == 0


def test_node_stage_parameter_presets_match_reference_defaults():
    g = A.excavation_params()          # excavated_surface_generator.cpp:28-47
    assert (g.depth, g.slope_angle_deg, g.offset_x, g.offset_y, g.point_density, g.terrain_search_radius) == (1.0, 75.0, 4.0, 1.0, 0.05, 0.5)
    assert (g.arm1_length, g.arm1_width, g.arm2_length, g.arm2_width, g.width, g.length) == (2.0, 1.2, 2.0, 1.2, 1.2, 1.8)
    assert g.l_shape_enabled == 1 and g.area_walls_vertical == 0
    d = A.DrivableParams()             # calc_drivable_area.cpp:21-26
    A.lib().plc_drivable_params_default(C.byref(d))
    assert (d.grid_resolution, d.map_width, d.map_height, d.max_gradient, d.min_points_per_cell, d.start_clear_radius) == \
           (1.0, 100.0, 100.0, 0.3, 10, 3.0)
    assert C.sizeof(A.ExcavationParams) == 12 * 8 + 8 and C.sizeof(A.DrivableParams) == 5 * 8 + 8
    assert C.sizeof(A.ExcavationInfo) == 4 * 8 + 3 * 8 and C.sizeof(A.DeviceTimes) == 12 * 4


def test_generator_sizing_function():
    # sizes depend on the parameters only (generateExcavatedSurface :487-584, generateExcavationArea :350-455)
    assert A.excavation_counts(A.excavation_params()) == (3115, 5575)                 # L: 1 680 bottoms + 205 edge points x 7 / x 19
    assert A.excavation_counts(A.excavation_params(l_shape_enabled=0)) == (1714, 3130)
    ns, na = A.excavation_counts(A.excavation_params(point_density=0.1))
    assert 0 < ns < 3115 and 0 < na < 5575
    for bad in (dict(point_density=0.0), dict(depth=-1.0), dict(slope_angle_deg=float("nan")), dict(slope_angle_deg=120.0),
                dict(point_density=1e-6),
                dict(point_density=0.001, slope_angle_deg=0.01)):        # 3 M entries x 5.7 M slope steps: not walkable
        with pytest.raises(A.PlcError):
            A.excavation_counts(A.excavation_params(**bad))
