#!/usr/bin/env python
"""Run `-m gpu` test files on the host emulator (tests/emu) instead of a device -- a developer's smoke run
of device-side changes when no GPU is at hand.  Slow (about 2e4 cell-subcycles per second): pick small cases
with -k.  Not a substitute for the hardware run and not part of any test tier.

    python tools/gpu_tests_on_emulator.py tests/test_gpu_parity.py -k "ragged or golden"
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import pytest  # noqa: E402

from tests.emu import emu  # noqa: E402

if __name__ == "__main__":
    with emu.patched(kernel=os.environ.get("EVP_B200_KERNEL", "tma")):
        sys.exit(pytest.main(["-m", "gpu", "-q", "--runxfail", "-p", "no:cacheprovider", *sys.argv[1:]]))
