#!/usr/bin/env python
"""The device sources under AddressSanitizer: the host build (tests/emu) compiled with -fsanitize=address and
driven through the C ABI on small tripole / displaced-pole cases, 1-4 ranks, both subcycle kernels.  Catches
accesses outside the arena, outside the dynamic shared-memory ring or the static shared variables, and outside
the caller's host arrays (not overruns from one plane into the next: the arena is one allocation).

    python tools/asan_on_emulator.py          # re-executes itself with libasan preloaded
"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
LIB = "/tmp/libevp_emu_asan.so"


def build():
    from tests.emu import emu
    emu.build()
    src = os.path.join(emu.BUILD, "src")
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address", "-fno-omit-frame-pointer", "-ffp-contract=off", "-fPIC",
           "-shared", "-pthread", "-DEVP_STRICT=1", "-w", "-I", emu.HERE, "-I", src, "-o", LIB,
           os.path.join(src, "evp_abi.cu.cpp"), os.path.join(emu.HERE, "cuda_emu.cpp")]
    subprocess.run(cmd, check=True)


def run():
    from cice5_b200 import evp, fixtures as fx
    from oracle import oracle as orc
    from tests.test_gpu_parity import OUT_CHECK
    from tests.util import clone_case, compare_cases
    lib = C.CDLL(LIB)
    lib.evp_gpu_strerror.restype = C.c_char_p
    lib.evp_gpu_last_cuda_error.restype = C.c_char_p
    evp._libs.update({True: lib, False: lib})
    for kernel in ("tma", "sweep"):
        os.environ["EVP_B200_KERNEL"] = kernel
        for g, scen, nr in ((fx.synthetic_tripole(96, 40, wet_poles=True), "full", 1), (fx.synthetic_tripole(30, 9), "full", 1),
                            (fx.synthetic_displaced(37, 11), "caps", 1), (fx.synthetic_tripole(96, 40), "full", 4),
                            (fx.synthetic_tripole(24, 12), "full", 2)):
            case = fx.make_case(g, scen, ndte=3, dt=3600.0)
            ref = clone_case(case)
            orc.run_case(ref)
            got = clone_case(case)
            if nr == 1:
                evp.run_case(got, strict=True)
            else:
                grp = evp.EvpGroup(case.dom, case.ndte, case.dt, nr, strict=True)
                try:
                    grp.set_grid(g.f)
                    grp.evp(case.dt, got.inp, got.io, got.out)
                finally:
                    grp.close()
            compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
            print(kernel, g.dom.name, "ranks", nr, "clean", flush=True)


if __name__ == "__main__":
    if os.environ.get("EVP_ASAN_CHILD") == "1":
        run()
    else:
        build()
        asan = subprocess.run(["gcc", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
        env = dict(os.environ, LD_PRELOAD=asan, EVP_ASAN_CHILD="1",
                   ASAN_OPTIONS="detect_leaks=0:detect_stack_use_after_return=0:halt_on_error=1")
        sys.exit(subprocess.run([sys.executable, os.path.abspath(__file__)], env=env).returncode)
