import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib():
    """libgraphslam_cuda.so, built in-tree if stale (nvcc cross-compiles without a GPU)."""
    from graphslam_b200 import _lib, build
    build.build()
    return _lib.load()


@pytest.fixture(scope="session")
def dbg():
    """libgraphslam_cuda_debug.so: the test-only CPU emulator and host hooks (never part of the product library)."""
    from graphslam_b200 import _lib, build
    build.build()
    return _lib.load_debug()
