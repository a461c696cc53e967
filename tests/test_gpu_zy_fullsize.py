"""Parity at the BASELINE sizes (configs[3] 1440x1080 and configs[4] 3600x2700), where the work
list has thousands of items, 12 / 30 strips, compact metric rows on almost every row segment and
the tripole fold runs at full width.  The property checked is the same as everywhere else: every inout
/ out field BITWISE equal to the CPU oracle -- including the configuration bench.py times: 0.1 degree,
ndte = 120, "full" and "caps" (the oracle needs about 20 s per call on the box's host cores).  The
3-rank group on one device runs a short call (it drives the launch-per-subcycle path)."""
import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.util import clone_case, compare_cases, phys
from tests.test_gpu_parity import OUT_CHECK, _oracle

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(900)]


def test_quarter_degree_one_call_bitwise():
    case = fx.make_case("tp025", "caps", ndte=120)
    ref = _oracle(case)
    got = clone_case(case)
    st = evp.run_case(got, strict=True, use_cuda_graph=True)
    assert st.active_t_cells > 100000
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    assert np.array_equal(phys(got.io["iceumask"]), phys(ref.io["iceumask"]))


@pytest.mark.parametrize("scenario,active", [("full", 7630364), ("caps", None)])
def test_tenth_degree_full_call_bitwise(scenario, active):
    """BASELINE configs[4] exactly as benchmarked: 3600 x 2700, ndte = 120, one whole evp call."""
    grid = fx.make_grid("tp01")
    case = fx.make_case(grid, scenario, ndte=120)
    ref = _oracle(case)
    got = clone_case(case)
    st = evp.run_case(got, strict=True, use_cuda_graph=True)
    if active is not None:
        assert st.active_t_cells == active
    assert st.active_t_cells > 1000000 and float(np.abs(got.io["uvel"]).max()) > 1.0e-3
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    assert np.array_equal(phys(got.io["iceumask"]), phys(ref.io["iceumask"]))


def test_tenth_degree_short_call_bitwise():
    grid = fx.make_grid("tp01")
    case = fx.make_case(grid, "full", ndte=6)
    ref = _oracle(case)
    got = clone_case(case)
    st = evp.run_case(got, strict=True)
    assert st.active_t_cells == 7630364
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    del got
    # three unequal x-slabs (strip-row balanced) on one device: same bits
    starts = fx.balanced_slab_starts(grid.f["tmask"][1:-1, 1:-1] > 0, 3)
    case4 = fx.make_case(grid, "full", ndte=4)
    ref4 = _oracle(case4)
    got4 = clone_case(case4)
    g = evp.EvpGroup(case4.dom, case4.ndte, case4.dt, 3, strict=True, slab_start=starts)
    try:
        g.set_grid(grid.f)
        g.evp(case4.dt, got4.inp, got4.io, got4.out)
    finally:
        g.close()
    compare_cases(got4, ref4, fields_out=OUT_CHECK, bitwise=True)
