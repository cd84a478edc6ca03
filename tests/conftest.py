import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU parity oracle (test infrastructure; never used by the product)."""
    from oracle import oracle as o
    o.build()
    return o


@pytest.fixture(scope="session")
def lib():
    """libdusp_b200.so, built in-tree if missing (nvcc cross-compiles without a GPU)."""
    from dusp_b200 import capi
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return capi.load()


@pytest.fixture(autouse=True)
def _seed():
    import dusp_b200
    dusp_b200.config.seed(12345)
    yield
