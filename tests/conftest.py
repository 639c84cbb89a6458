import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def bre():
    """ctypes binding module of libbre.so (builds the library if it is missing)."""
    mod = ge.load_bre()
    if not os.path.exists(mod.LIB_PATH):
        ge.build()
    return mod


@pytest.fixture(scope="session")
def orc():
    """ctypes binding module of the CPU oracle (test infrastructure)."""
    return ge.load_orc()


@pytest.fixture()
def oracle(orc):
    return orc.Oracle()


@pytest.fixture()
def ctx(bre):
    c = bre.Bre(device=0)
    yield c
    c.close()
