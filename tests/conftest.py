import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))
if HERE not in sys.path:
    sys.path.insert(0, HERE)   # tests/textio.py, tests/evalpath_text.py


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.binding import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def refshim():
    from oracle.binding import RefShim
    try:
        return RefShim()
    except FileNotFoundError:
        pytest.skip("oracle/_ref not built (reference sources absent)")


@pytest.fixture(scope="session")
def ref_binary():
    from oracle.binding import ref_binary
    p = ref_binary()
    if p is None:
        pytest.skip("oracle/_ref/gfalign not built (reference sources absent)")
    return p


@pytest.fixture(scope="session")
def engine():
    import gfalign_b200
    eng = gfalign_b200.Engine(0)   # raises without a GPU: gpu tests must not pass on a fallback
    yield eng
    eng.close()
