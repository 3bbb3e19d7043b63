import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as g

    return g.load_package()


@pytest.fixture(scope="session")
def ref():
    """The reference oracle (oracle/_ref/libsgref.so = unmodified reference headers)."""
    import oracle_ref

    r = oracle_ref.load()
    if r is None:
        pytest.skip("oracle/_ref/libsgref.so not built (reference not mounted)")
    return r
