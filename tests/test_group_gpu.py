"""plc_group_*: one process driving several GPUs.  A group of one device runs everywhere;
the two-device cases need a box with at least two GPUs (`gpurun --gpus 2`) and are skipped otherwise."""
import ctypes as C

import numpy as np
import pytest

from pointcloud_processor_b200 import Context, _abi as A, scenes
from pointcloud_processor_b200.context import Group

pytestmark = pytest.mark.gpu


def device_count() -> int:
    n = C.c_int(0)
    rt = C.CDLL("libcudart.so.12")
    return n.value if rt.cudaGetDeviceCount(C.byref(n)) == 0 else 0


def single(sc, observed=None, **kw):
    with Context(0) as ctx:
        ctx.set_terrain(sc.terrain)
        ctx.set_cells(sc.cells)
        if observed is not None:
            ctx.set_observed(observed)
        return ctx.sweep(sc.params, sc.fixed, sc.candidates, **kw)


def same(a, b):
    assert a.best_index == b.best_index and a.best_score == b.best_score and a.best_covered == b.best_covered
    assert np.array_equal(a.scores, b.scores) and np.array_equal(a.covered, b.covered)
    assert np.array_equal(a.mask, b.mask)
    assert np.array_equal(a.cell_flags, b.cell_flags)
    assert np.array_equal(a.fixed_scores, b.fixed_scores) and a.fixed_total == b.fixed_total
    assert a.counters == b.counters


@pytest.mark.parametrize("n_dev", [1, 2, 4])
def test_group_equals_single_context(n_dev):
    if device_count() < n_dev:
        pytest.skip(f"needs {n_dev} GPUs")
    sc = scenes.build("small")
    observed = sc.area[::9].copy()
    ref = single(sc, observed, mask=True, flags=True, fixed_scores=True)
    with Group(list(range(n_dev))) as g:
        g.set_terrain(sc.terrain)
        g.set_observed(observed)
        g.set_cells(sc.cells)
        got = g.sweep(sc.params, sc.fixed, sc.candidates, mask=True, flags=True, fixed_scores=True)
        same(got, ref)
        # fewer candidates than devices, and none at all
        few = g.sweep(sc.params, sc.fixed, sc.candidates[:1], mask=True)
        assert few.best_index == 0 and few.scores[0] == ref.scores[0]
        none = g.sweep(sc.params, sc.fixed, sc.candidates[:0])
        assert none.best_index == -1 and none.best_score == -np.inf


def test_group_rejects_bad_device_lists():
    L = A.lib()
    h = C.c_void_p()
    devs = (C.c_int * 2)(0, 0)
    assert L.plc_group_create(devs, 2, C.byref(h)) == A.PLC_E_INVALID_ARG
    assert L.plc_group_create(devs, 0, C.byref(h)) == A.PLC_E_INVALID_ARG
    bad = (C.c_int * 1)(4096)
    assert L.plc_group_create(bad, 1, C.byref(h)) != A.PLC_OK


@pytest.mark.parametrize("n_dev", [1, 2])
def test_node_on_group_matches_node_on_context(n_dev):
    if device_count() < n_dev:
        pytest.skip(f"needs {n_dev} GPUs")
    from pointcloud_processor_b200.nodes import VirtualLidarNode
    from test_nodes_gpu import xyzrgb_records
    terrain = scenes.terrain("L", 40_000, seed=21, side=24.0)
    area, _ = scenes.excavation_area("L", 0.05)

    def cycle(owner):
        node = VirtualLidarNode(owner)
        node.set_parameter("num_candidates", 81)
        node.excavation_area_callback(xyzrgb_records(area))
        node.terrain_callback(xyzrgb_records(terrain))
        node.set_base_transform((0.0, 0.0, 0.0))
        first = node.run_optimization()
        node.set_parameter("max_distance", 7.0)
        node.zx120_points_callback(xyzrgb_records((area[::7] + np.float32(0.03)).astype(np.float32)))
        second = node.run_optimization()
        out = [(r.ran, r.n_candidates, r.best.x, r.best.y, r.best.z, r.best_score, r.fixed_total, r.visibility_ok)
               for r in (first, second)]
        flags = node.cells()[1].copy()
        node.close()
        return out, flags

    with Context(0) as ctx:
        want, want_flags = cycle(ctx)
    with Group(list(range(n_dev))) as g:
        got, got_flags = cycle(g)
    assert got == want
    assert np.array_equal(got_flags, want_flags)
