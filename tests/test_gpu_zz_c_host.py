"""The C-ABI boundary from plain C (examples/evp_host.c): no Python, no torch in the loop.

CPU part (runs everywhere): the example compiles against include/evp_b200.h, links the in-tree
library and, without a device, fails loudly with EVP_ERR_NO_DEVICE (there is no CPU fallback).
GPU part: the same binary runs evp(dt) on a closed basin and must reproduce the oracle's numbers
for that case (computed with oracle.run_case on the same inputs: uvel = 1.154007e-01,
vvel = -2.511930e-02 at the centre).  Validated on hardware since the end of round 1
(no xfail mark: a failure is a failure)."""
import os
import re
import shutil
import subprocess

import pytest

from cice5_b200 import build as _build, evp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _compile(tmp_path):
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    evp.load_library(True)                                  # builds the library if it is stale
    libdir = os.path.dirname(_build.lib_path(True))
    exe = str(tmp_path / "evp_host")
    r = subprocess.run([cc, "-O2", "-std=c99", "-Wall", "-I", os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "examples", "evp_host.c"), "-L", libdir, "-levp_b200_strict",
                        f"-Wl,-rpath,{libdir}", "-lm", "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_c_host_links_and_fails_loudly_without_a_device(tmp_path, have_gpu):
    if have_gpu:
        pytest.skip("a GPU is present")
    r = subprocess.run([_compile(tmp_path)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 5 and "no CPU fallback" in r.stderr          # EVP_ERR_NO_DEVICE


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_c_host_runs_evp_and_matches_the_oracle(tmp_path):
    r = subprocess.run([_compile(tmp_path)], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"uvel = (\S+)\s+vvel = (\S+)", r.stdout)
    u, v = float(m.group(1)), float(m.group(2))
    assert abs(u - 1.154007e-01) <= 2e-6 * 1.154007e-01 and abs(v + 2.511930e-02) <= 2e-6 * 2.511930e-02
    assert "5828 active T cells" in r.stdout
