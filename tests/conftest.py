import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def engine():
    """One Engine on cuda:0 shared by the GPU tests (fails loudly without the CUDA library)."""
    from mpmc_b200 import Engine

    e = Engine(0)
    yield e
    e.close()
