import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gaussian-splatting-slam_b200")
for p in (ROOT, PKG, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def gsr_lib():
    """Builds (if stale) and loads libgsr.so -- needs nvcc, not a GPU."""
    sys.path.insert(0, PKG)
    import build as gsr_build
    gsr_build.build()
    from diff_gaussian_rasterization import _C
    return _C.lib()
