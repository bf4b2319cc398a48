"""pytest configuration.  `-m "not gpu"` runs here (no GPU): oracle, host logic, ABI surface.
`-m gpu` runs on a B200: parity of the CUDA path against the oracle, through the C ABI."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if os.fspath(ROOT) not in sys.path:
    sys.path.insert(0, os.fspath(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def built_libraries():
    """Both shared libraries are built in-tree by __graft_entry__.build(); rebuild only if absent."""
    from pointcloud_processor_b200 import _abi
    if not _abi.LIB_PATH.exists():
        subprocess.check_call(["make", "-C", os.fspath(ROOT / "pointcloud_processor_b200" / "csrc")])
    from oracle import binding
    binding.build()
    return True


@pytest.fixture(scope="session")
def scene_cache():
    from pointcloud_processor_b200 import scenes
    cache = {}

    def get(name, **kw):
        key = (name, tuple(sorted(kw.items())))
        if key not in cache:
            cache[key] = scenes.build(name, **kw)
        return cache[key]
    return get


def assert_same_outputs(fast, got):
    """The any-hit traversal (the library's default) against the counting traversal of the same sweep: every output
    the reference has, bit for bit; only the evaluated-sample counter is not produced."""
    import numpy as np
    assert (fast.best_index, fast.best_score, fast.best_covered, fast.fixed_total) == \
           (got.best_index, got.best_score, got.best_covered, got.fixed_total)
    for name in ("scores", "covered", "mask", "cell_flags", "fixed_scores"):
        a, b = getattr(fast, name), getattr(got, name)
        assert (a is None) == (b is None), name
        if a is not None:
            assert np.array_equal(a, b, equal_nan=True), name
    for k in ("rays", "rays_gated_in", "rays_visible", "fixed_samples"):
        assert fast.counters[k] == got.counters[k], k
    assert fast.counters["samples"] == 0


@pytest.fixture(scope="module")
def ctx():
    """A context in counting mode (the tests compare the evaluated-sample counter with the oracle's) whose every
    sweep is repeated in the default any-hit mode and required to give identical outputs."""
    from pointcloud_processor_b200 import Context
    c = Context(0)
    c.count_samples(True)

    def both(plain):
        def run(*a, **kw):
            got = plain(*a, **kw)
            if c.counting:
                c.count_samples(False)
                try:
                    fast = plain(*a, **kw)
                finally:
                    c.count_samples(True)
                assert_same_outputs(fast, got)
            return got
        return run
    c.sweep = both(c.sweep)
    c.sweep_strided = both(c.sweep_strided)
    yield c
    c.close()
