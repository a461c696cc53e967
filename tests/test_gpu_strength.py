"""ice_strength on the device against the CPU oracle.

reference: source/ice_mechred.F90:2111-2269 (ice_strength), :936-1285 (ridge_itd); evp calls it on
the icellt list (source/ice_dyn_evp.F90:291-301).

Parity bar: BITWISE in every mode.  The only transcendental on the path, exp (exponential participation
function, the default; Hibler's kstrength = 0), is the correctly rounded evp_exp of csrc/evp_exp.h on both
sides (oracle: cr_exp=True).  Round 1 compared CUDA's exp with libm's to a few ulp and the gx3 "caps" state
failed its 1e-10 state bar on a B200: one ulp of `strength` is amplified by 120 subcycles (DESIGN.md
section 3)."""
import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.util import STATE, REL_TOL, clone_case, compare_cases, phys

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def _run_pair(grid, scenario, ndte, **modes):
    from oracle import oracle as orc
    g = fx.make_grid(grid)
    case = fx.make_case(g, scenario, ndte=ndte)
    itd = fx.ice_itd(case.inp)
    # oracle: strength on every cell, evp keeps it on the icellt list (strength_masked = 1)
    ref = clone_case(case)
    ref.inp["strength"] = orc.ice_strength(g.dom, case.inp["aice"], case.inp["vice"], itd["aice0"],
                                           itd["aicen"], itd["vicen"], cr_exp=True, **modes)
    orc.run_case(ref, strength_masked=0 if modes.get("kstrength", 1) == 0 else 1)
    got = clone_case(case)
    got.inp.pop("strength")                       # the device computes it
    e = evp.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    try:
        e.set_grid(g.f)
        e.upload_itd(itd["aicen"], itd["vicen"], itd["aice0"], **modes)
        e.evp(case.dt, got.inp, got.io, got.out)
    finally:
        e.close()
    return got, ref


@pytest.mark.parametrize("krdg_redist", [0, 1])
def test_strength_thorndike_participation_bitwise(krdg_redist):
    got, ref = _run_pair("tp1", "caps", 30, krdg_partic=0, krdg_redist=krdg_redist)
    assert np.array_equal(phys(got.out["strength_out"]), phys(ref.out["strength_out"]))
    compare_cases(got, ref, fields_io=STATE, fields_out=("divu", "shear", "strocnxT"), bitwise=True)


@pytest.mark.parametrize("grid,scenario", [("gx3", "full"), ("gx3", "caps"), ("tp1", "caps")])
def test_strength_default_modes_bitwise(grid, scenario):
    """krdg_partic = 1, krdg_redist = 1 (the namelist defaults): exp on both sides is evp_exp."""
    got, ref = _run_pair(grid, scenario, 120)
    a, b = phys(got.out["strength_out"]), phys(ref.out["strength_out"])
    assert (b > 0).any()
    assert np.array_equal(a, b), f"strength differs in {np.count_nonzero(a != b)} cells, max rel " \
                                 f"{np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)):.3e}"
    compare_cases(got, ref, fields_io=STATE, fields_out=("divu", "shear", "strocnxT"), bitwise=True)


def test_strength_hibler_bitwise():
    """kstrength = 0: Pstar*vice*exp(-Cstar*(1-aice)) on the physical cells (ice_mechred.F90:2259-2265)."""
    got, ref = _run_pair("gx3", "caps", 30, kstrength=0)
    assert np.array_equal(phys(got.out["strength_out"]), phys(ref.out["strength_out"]))
    compare_cases(got, ref, fields_io=STATE, fields_out=("divu", "shear"), bitwise=True)


def test_principal_stress_from_resident_state_bitwise():
    """sig1 / sig2 (ice_dyn_shared.F90:853-893) from the device-resident stresses and prs_sig of
    the last evp call == the oracle's principal_stress on the downloaded fields, bit for bit."""
    from oracle import oracle as orc
    g = fx.make_grid("gx3")
    case = fx.make_case(g, "full", ndte=120)
    got = clone_case(case)
    e = evp.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    try:
        e.set_grid(g.f)
        e.evp(case.dt, got.inp, got.io, got.out)
        s1, s2 = e.principal_stress()
    finally:
        e.close()
    r1, r2 = orc.principal_stress(got.io["stressp_1"], got.io["stressm_1"], got.io["stress12_1"],
                                  got.out["prs_sig"])
    assert np.array_equal(s1, r1) and np.array_equal(s2, r2)
    assert (s1 != 1.0e30).any() and (s1 == 1.0e30).any()
