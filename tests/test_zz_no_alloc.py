"""evp_gpu_principal_stress_static: the allocation-free entry point (no counterpart in the reference)
against the existing evp_gpu_principal_stress, through the C ABI, bit for bit (the
tolerance of every device test here), on two calls in a row with different data -- the static scratch keeps the last
call's contents and nothing of it may leak into the next result.  Collected last on purpose (file name).
The body also runs on the host build of the device sources in the CPU tier."""
import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.util import clone_case


def _check_static_against_handle(dev):
    # two different states: gx3 (102 x 118 block) after one evp call, then after a second call with other forcing
    g = fx.make_grid("gx3")
    case = fx.make_case(g, "full", ndte=20)
    got = clone_case(case)
    e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    results = []
    try:
        e.set_grid(g.f)
        for call in range(2):
            if call == 1:
                got.inp.update({k: v.copy() for k, v in fx.forcing(g, case.inp, 60, steps_per_cycle=240).items()})
            e.evp(case.dt, got.inp, got.io, got.out)
            s1, s2 = e.principal_stress()                             # existing entry point (arena + staging scratch)
            t1, t2 = dev.principal_stress_static(got.io["stressp_1"], got.io["stressm_1"], got.io["stress12_1"],
                                                 got.out["prs_sig"])   # new entry point (static device scratch)
            assert np.array_equal(t1, s1) and np.array_equal(t2, s2), f"call {call}"
            assert (s1 != 1.0e30).any() and (s1 == 1.0e30).any()
            results.append(t1.copy())
    finally:
        e.close()
    assert not np.array_equal(results[0], results[1])                 # the two calls really saw different data
    # a smaller block right after a larger one: stale scratch beyond the new size must not matter
    rng = np.random.default_rng(3)
    small = [np.ascontiguousarray(rng.standard_normal((9, 7))) for _ in range(3)] + [np.abs(rng.standard_normal((9, 7)))]
    t1, t2 = dev.principal_stress_static(*small)
    root = np.sqrt(small[1] * small[1] + 4.0 * (small[2] * small[2]))
    assert np.array_equal(t1, (0.5 * (small[0] + root)) / small[3])
    assert np.array_equal(t2, (0.5 * (small[0] - root)) / small[3])
    # beyond the stated maximum: refused with a clear error, nothing allocated
    n = dev.STATIC_MAX_CELLS + 1
    big = np.zeros((1, n))
    with pytest.raises(dev.EvpError, match="static scratch"):
        dev.principal_stress_static(big, big, big, big)


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_principal_stress_static_matches_existing_entry_point_twice():
    _check_static_against_handle(evp)


def test_principal_stress_static_on_the_host_build():
    from tests.emu import emu
    with emu.patched(kernel="tma") as dev:
        _check_static_against_handle(dev)
