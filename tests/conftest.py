import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib_built():
    """Make sure libsvtk_b200.so exists (nvcc cross-compiles here without a GPU)."""
    from svtrek_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def auditor(lib_built):
    from svtrek_b200 import api
    a = api.Auditor(0)
    yield a
    a.close()
