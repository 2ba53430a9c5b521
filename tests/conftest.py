import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_py
    return oracle_py.load()


@pytest.fixture(scope="session")
def ctx():
    """One cg_ctx on cuda:0 through the C-ABI; fails loudly when the library or GPU is missing."""
    from conga_b200 import Context
    c = Context(0)
    yield c
    c.close()
