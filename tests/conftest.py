import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))   # oracle_lib: the checker (test infrastructure)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pgpu():
    """The product package (CUDA path).  Fails loudly when the library or a GPU is missing."""
    import presto_b200
    return presto_b200


@pytest.fixture(scope="session")
def ctx(pgpu):
    return pgpu.default_context()
