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
    """the product package (directory name has a hyphen, so it is imported by name).  A checkout without build outputs (the .so is
    git-ignored) is built here first where nvcc exists: the tests need the library, and the product itself never builds or falls
    back on its own"""
    pkg = importlib.import_module("pbrt-minus_b200")
    if not os.path.exists(pkg.api.LIB_PATH) and os.path.exists(os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")):
        sys.path.insert(0, os.path.join(ROOT, "pbrt-minus_b200"))
        import build as _build
        _build.build()
    return pkg


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle


def has_gpu():
    """True when a CUDA device is there.  Only "no device" turns the GPU suite into skips: a missing or broken
    libb200pt.so (nvcc failure, undefined symbol) is an error wherever a device exists, and everywhere when
    B200PT_REQUIRE_GPU=1 (set it on the GPU box)."""
    require = os.environ.get("B200PT_REQUIRE_GPU") == "1"
    try:
        import torch
        device = torch.cuda.is_available()
    except Exception:
        device = os.path.exists("/dev/nvidia0")
    if not device and not require:
        return False
    m = importlib.import_module("pbrt-minus_b200")
    n = m.load_library().b200pt_device_count()     # raises if the library is not built / does not load
    if n <= 0 and require:
        raise RuntimeError("B200PT_REQUIRE_GPU=1 but libb200pt sees no CUDA device")
    return n > 0
