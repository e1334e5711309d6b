import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def b200pt():
    """the product package (directory name has a hyphen, so it is imported by name)"""
    return importlib.import_module("pbrt-minus_b200")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle


def has_gpu():
    try:
        m = importlib.import_module("pbrt-minus_b200")
        return m.load_library().b200pt_device_count() > 0
    except Exception:
        return False
