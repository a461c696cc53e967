"""nvcc driver: builds the in-tree C-ABI libraries for sm_100a.

  lib/libevp_b200_strict.so  PRODUCT build: -O3 -fmad=false, IEEE evaluation without contraction,
                             bitwise equal to the oracle (and to itself across decompositions)
  lib/libevp_b200.so         same sources with FMA contraction on (comparison only)
"""
from __future__ import annotations

import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIBDIR = os.path.join(_HERE, "lib")
SOURCES = ["evp_abi.cu"]
HEADERS = ["evp_kernels.cuh", "evp_prep.cuh", "evp_subcycle_tma.cuh", "evp_strength.cuh"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]

VARIANTS = {
    "libevp_b200.so": [],
    "libevp_b200_strict.so": ["-fmad=false", "-DEVP_STRICT=1"],
}


def lib_path(strict: bool = False) -> str:
    return os.path.join(LIBDIR, "libevp_b200_strict.so" if strict else "libevp_b200.so")


def _stale(out: str) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    deps.append(os.path.join(_HERE, "..", "include", "evp_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> list[str]:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    os.makedirs(LIBDIR, exist_ok=True)
    outs = []
    for name, extra in VARIANTS.items():
        out = os.path.join(LIBDIR, name)
        outs.append(out)
        if not force and not _stale(out):
            continue
        if not os.path.exists(nvcc):
            raise RuntimeError(f"nvcc not found and {out} is missing or stale")
        cmd = [nvcc, *ARCH, "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-shared",
               *(["-Xptxas", "-v"] if verbose else []), *extra, "-o", out,
               *[os.path.join(CSRC, s) for s in SOURCES]]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose:
            print(r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    return outs


if __name__ == "__main__":
    import sys
    print("\n".join(build(force="--force" in sys.argv, verbose="-v" in sys.argv)))
