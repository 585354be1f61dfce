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
    from oracle.loader import Port
    return Port()


@pytest.fixture(scope="session")
def ref():
    """The untouched reference header (oracle/_ref); skipped if neither sources nor a prebuilt .so exist."""
    from oracle.loader import Ref
    try:
        return Ref()
    except (OSError, FileNotFoundError) as e:
        pytest.skip("reference shim unavailable: %s" % e)


@pytest.fixture(scope="session")
def oracle(request):
    """Best available checker: the reference itself when its shim loads, else the C restatement."""
    from oracle.loader import best
    return best()
