import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    import orc as _orc  # oracle/orc.py (test infrastructure)
    _orc.lib()
    return _orc


@pytest.fixture(scope="session")
def ctx():
    """One CUDA context for the whole GPU session; creating it fails loudly when there is no device."""
    from pyvspg import api
    c = api.Context(0)
    yield c
    c.close()
