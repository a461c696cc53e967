"""pytest configuration: the `gpu` marker (tests that need a real B200) and shared fixtures."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _have_gpu() -> bool:
    try:
        from cice5_b200 import evp
        return evp.device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without a device must fail loudly, not skip: there is no CPU fallback.
    pass


@pytest.fixture(scope="session")
def have_gpu():
    return _have_gpu()


@pytest.fixture(params=["tma", "sweep"])
def subcycle_kernel(request, monkeypatch):
    """Run a test once per subcycle-kernel variant (v2 TMA ring = product default, v1 register
    sweep); the library reads EVP_B200_KERNEL when a handle is created."""
    monkeypatch.setenv("EVP_B200_KERNEL", request.param)
    return request.param
