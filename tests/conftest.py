import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def lib():
    """The in-tree shared library; built on demand here (nvcc cross-compiles without a GPU)."""
    from theroundtaible_b200.build import build
    build()
    from theroundtaible_b200 import engine
    return engine.load_library()
