"""Revised EVP (Bouillon et al. 2013) on the device against the CPU oracle.

reference: set_evp_parameters `revised_evp` branch (source/ice_dyn_shared.F90:225-233), stresses
restarted from zero in evp_prep2 (:492-504), revp in stepu (:715,732-734).

The CPU oracle side is tested in tests/test_oracle_cpu.py.  Validated on hardware since the end of round 1
(no xfail mark: a failure is a failure)."""
import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.util import clone_case, compare_cases, phys
from tests.test_gpu_parity import OUT_CHECK

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300), pytest.mark.usefixtures("subcycle_kernel")]


def _xmin(grid):
    tm = grid.f["tmask"][1:-1, 1:-1] > 0
    return float(min(grid.f["dxt"][1:-1, 1:-1][tm].min(), grid.f["dyt"][1:-1, 1:-1][tm].min()))


@pytest.mark.parametrize("grid,scenario", [("gx3", "full"), ("tp1", "caps")])
def test_revised_evp_one_call_bitwise(grid, scenario):
    from oracle import oracle as orc
    case = fx.make_case(grid, scenario, ndte=120)
    ref = clone_case(case)
    orc.run_case(ref, revised_evp=True)
    got = clone_case(case)
    evp.run_case(got, strict=True, revised_evp=True)          # xmin taken from the block
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    assert np.array_equal(phys(got.io["iceumask"]), phys(ref.io["iceumask"]))


def test_revised_evp_two_calls_and_four_ranks_bitwise():
    from oracle import oracle as orc
    grid = fx.synthetic_tripole(96, 40)
    case = fx.make_case(grid, "full", ndte=12, dt=3600.0)
    ref = clone_case(case)
    got = clone_case(case)
    g = evp.EvpGroup(case.dom, case.ndte, case.dt, 4, strict=True, revised_evp=True, xmin=_xmin(grid))
    try:
        g.set_grid(grid.f)
        for _ in range(2):
            ref.out = fx.zeroed_outputs(case.dom)
            got.out = fx.zeroed_outputs(case.dom)
            orc.run_case(ref, revised_evp=True)
            g.evp(case.dt, got.inp, got.io, got.out)
            compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    finally:
        g.close()
