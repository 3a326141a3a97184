import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _have_gpu():
    try:
        from igm_b200 import _abi
        return _abi.load().igmk_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def _built():
    """The native pieces must exist before anything is tested (build is a no-op when up to date)."""
    import __graft_entry__
    __graft_entry__.build()


@pytest.fixture(scope="session")
def ff_golden():
    return np.load(os.path.join(GOLDEN, "ff_golden.npz"))


@pytest.fixture(scope="session")
def astep_golden():
    return np.load(os.path.join(GOLDEN, "astep_golden.npz"))
