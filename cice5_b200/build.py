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
HEADERS = ["evp_kernels.cuh", "evp_prep.cuh", "evp_subcycle_tma.cuh", "evp_loop_tma.cuh", "evp_strength.cuh", "evp_remap.cuh", "evp_eap.cuh", "evp_exp.h", "evp_trig.h"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]

VARIANTS = {
    "libevp_b200.so": [],
    "libevp_b200_strict.so": ["-fmad=false", "-DEVP_STRICT=1"],
}


class NoCompiler(RuntimeError):
    """nvcc is not on this box (the GPU box runs the prebuilt libraries)."""


def lib_path(strict: bool = False) -> str:
    return os.path.join(LIBDIR, "libevp_b200_strict.so" if strict else "libevp_b200.so")


def _stale(out: str) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    deps.append(os.path.join(_HERE, "..", "include", "evp_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def is_stale(out: str) -> bool:
    return _stale(out)


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
            raise NoCompiler(f"nvcc not found and {out} is missing or stale")
        cmd = [nvcc, *ARCH, "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-shared",
               *(["-Xptxas", "-v"] if verbose else []), *extra, "-o", out,
               *[os.path.join(CSRC, s) for s in SOURCES]]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose:
            print(r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    return outs


def build_variant(name: str, defines: list[str], strict: bool = True) -> str:
    """Experiment build (kernel-tuning knobs, diagnosis options): lib/exp_<name>.so from the same
    sources with extra -D flags; load it with EVP_B200_LIB=<path>.  Never the product library."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    os.makedirs(LIBDIR, exist_ok=True)
    out = os.path.join(LIBDIR, f"exp_{name}.so")
    extra = VARIANTS["libevp_b200_strict.so" if strict else "libevp_b200.so"]
    cmd = [nvcc, *ARCH, "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-shared", *extra, *defines,
           "-o", out, *[os.path.join(CSRC, s) for s in SOURCES]]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    return out


if __name__ == "__main__":
    import sys
    if "--variant" in sys.argv:        # python -m cice5_b200.build --variant mergeuv -DEVP_TMA_MERGE_UV=1
        i = sys.argv.index("--variant")
        print(build_variant(sys.argv[i + 1], [a for a in sys.argv[i + 2:] if a.startswith("-D")]))
    else:
        print("\n".join(build(force="--force" in sys.argv, verbose="-v" in sys.argv)))
