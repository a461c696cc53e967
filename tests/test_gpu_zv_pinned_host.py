"""evp_gpu_pin_host_arrays on hardware: pageable numpy arrays page-locked in place, same answers, and the
host-buffer call is not slower than with pageable arrays.

reference: the arrays that cross the boundary are static module arrays (source/ice_state.F90:54-56,143-146;
source/ice_flux.F90:37-107), so a Fortran host can register them once.

Validated on hardware since the end of round 1 (no xfail mark: a failure is a failure)."""
import time

import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.util import clone_case, compare_cases
from tests.test_gpu_parity import OUT_CHECK

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]


def _timed_calls(e, case, n=3):
    e.evp(case.dt, case.inp, case.io, case.out)
    t0 = time.perf_counter()
    for _ in range(n):
        e.evp(case.dt, case.inp, case.io, case.out)
    return (time.perf_counter() - t0) / n


def test_pinned_host_arrays_same_answers_and_not_slower():
    from oracle import oracle as orc
    grid = fx.make_grid("tp025")
    case = fx.make_case(grid, "full", ndte=4, dt=fx.GRID_DT.get("tp025", 3600.0))
    ref = clone_case(case)
    orc.run_case(ref)
    got = clone_case(case)
    e = evp.EvpGpu.for_domain(got.dom, got.ndte, got.dt, strict=True)
    try:
        e.set_grid(grid.f)
        e.evp(got.dt, got.inp, got.io, got.out)
        compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
        pageable = _timed_calls(e, clone_case(case))
        pinned_case = clone_case(case)
        e.pin_host(pinned_case.inp, pinned_case.io, pinned_case.out)
        e.pin_host(pinned_case.inp, pinned_case.io, pinned_case.out)        # twice is harmless
        pinned = _timed_calls(e, pinned_case)
        again = clone_case(case)
        e.pin_host(again.inp, again.io, again.out)
        e.evp(again.dt, again.inp, again.io, again.out)
        compare_cases(again, ref, fields_out=OUT_CHECK, bitwise=True)
    finally:
        e.close()
    print(f"evp_gpu_run per call: pageable {pageable * 1e3:.1f} ms, page-locked {pinned * 1e3:.1f} ms")
    assert pinned < 1.1 * pageable
