import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the C oracle and the CUDA library exist (both build without a GPU)."""
    from oracle import pyoracle
    pyoracle.build()
    from psgradientsdf_b200 import build as b
    if b.needs_build():
        b.build()
