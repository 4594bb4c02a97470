import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port():
    from oracle import oracle
    return oracle.Port()


@pytest.fixture(scope="session")
def ref():
    from oracle import oracle
    if not oracle.Ref.available():
        pytest.skip("oracle/_ref not built here (needs /root/reference at build time)")
    return oracle.Ref()


@pytest.fixture(scope="session")
def checker():
    """Strongest CPU checker available: the reference's own headers, else the C port."""
    from oracle import oracle
    return oracle.best()
