import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "host_harness")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def built_library():
    """The in-tree C-ABI library, built on demand (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as ge
    return ge.build()


@pytest.fixture(scope="session")
def hostmath():
    """g++ build of the product's __host__ __device__ arithmetic (test infrastructure only)."""
    import ctypes
    import build as hb
    return ctypes.CDLL(hb.build())
