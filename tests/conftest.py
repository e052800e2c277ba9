import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def hv_lib():
    """The built C-ABI library (compiled on demand; nvcc cross-compiles without a GPU)."""
    from hive_b200 import build, _lib
    build.build()
    return _lib.load()


@pytest.fixture(scope="session")
def c_oracle():
    from oracle import c_oracle
    c_oracle.build()
    return c_oracle
