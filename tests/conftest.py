"""pytest configuration: the `gpu` marker (tests that need a real B200) and shared fixtures."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "isolated: device test run in a child pytest process of its own "
                                       "(code on its first hardware run); a failure of the child is a failure")


def _have_gpu() -> bool:
    try:
        from cice5_b200 import evp
        return evp.device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without a device must fail loudly, not skip: there is no CPU fallback.
    # A GPU test whose kernel never returns blocks inside a C call, where pytest-timeout's default
    # signal method cannot interrupt it: device tests without an explicit method get the thread
    # method (the run is ended with a stack dump instead of occupying the box until it is reclaimed).
    for item in items:
        if item.get_closest_marker("gpu") is None:
            continue
        tm = item.get_closest_marker("timeout")
        if tm is None:
            item.add_marker(pytest.mark.timeout(900, method="thread"))
        elif "method" not in tm.kwargs and len(tm.args) < 2:
            item.add_marker(pytest.mark.timeout(*tm.args, method="thread", **tm.kwargs), append=False)


def _pending_on_hardware(item) -> bool:
    return item.get_closest_marker("gpu") is not None and item.get_closest_marker("isolated") is not None


@pytest.hookimpl(tryfirst=True)
def pytest_pyfunc_call(pyfuncitem):
    """Device tests marked `isolated` (code on its first hardware run) execute in a child pytest
    process of their own, killed at its deadline: a hang or a sticky CUDA error in code that has
    never met a device cannot stall or poison the rest of the suite.  No xfail is involved: a
    failing child fails the test (round 1 hid a real failure behind xfail(strict=False))."""
    import subprocess
    if not _pending_on_hardware(pyfuncitem) or os.environ.get("EVP_TEST_CHILD"):
        return None
    tm = pyfuncitem.get_closest_marker("timeout")
    limit = max(30.0, float(tm.args[0]) - 30.0) if tm and tm.args else 270.0
    env = dict(os.environ, EVP_TEST_CHILD="1")
    cmd = [sys.executable, "-m", "pytest", pyfuncitem.nodeid, "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider"]
    try:
        r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=limit)
    except subprocess.TimeoutExpired:
        pytest.fail(f"child process of {pyfuncitem.nodeid} killed after {limit:.0f} s", pytrace=False)
    if r.returncode != 0:
        pytest.fail("child pytest failed:\n" + (r.stdout + r.stderr)[-4000:], pytrace=False)
    return True


@pytest.fixture(scope="session")
def have_gpu():
    return _have_gpu()


@pytest.fixture(params=["tma", "sweep"])
def subcycle_kernel(request, monkeypatch):
    """Run a test once per subcycle-kernel variant (v2 TMA ring = product default, v1 register
    sweep); the library reads EVP_B200_KERNEL when a handle is created."""
    monkeypatch.setenv("EVP_B200_KERNEL", request.param)
    return request.param
