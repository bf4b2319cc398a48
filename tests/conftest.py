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


@pytest.fixture(scope="module")
def ctx():
    from pointcloud_processor_b200 import Context
    c = Context(0)
    yield c
    c.close()
