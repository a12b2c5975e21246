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
def engine():
    """The product library through its ctypes binding (fails loudly if it was not built)."""
    import importlib

    return importlib.import_module("particle-simulation_b200")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib

    return oracle_lib
