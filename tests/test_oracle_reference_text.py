"""The oracle against vectors produced by the REFERENCE'S OWN SUBROUTINE TEXTS.

tests/golden/ref_block_vectors.npz was written by tests/golden/make_reference_vectors.py, which
executes set_evp_parameters / evp_prep1 / to_ugrid / evp_prep2 / stress / stepu / evp_finish /
to_tgrid / principal_stress / ice_strength (+ ridge_itd) mechanically from the reference checkout
(tests/golden/fortran_exec.py, a Fortran-subset interpreter) on seeded 12 x 10 blocks, halo updates
left out.  The C oracle in `no_halo` mode has to reproduce every output BIT FOR BIT, for the
default build, `-Dcoupled`, the AusCOM build with a non-zero ocean turning angle (with and without
use_ocnslope) and revised EVP.  Nothing here reads /root/reference.
"""
import os

import numpy as np
import pytest

from cice5_b200 import fixtures as fx
from oracle import oracle as orc

NPZ = os.path.join(os.path.dirname(__file__), "golden", "ref_block_vectors.npz")
_ANGLE = 25.0 * np.pi / 180.0
# oracle configuration of each variant (the generator's VARIANTS table, restated so that the test does
# not import the generator, which needs the reference checkout)
VARIANTS = {
    "default": {},
    "coupled": {"tilt_mode": 1},
    "auscom_slope": {"tilt_mode": 1, "cosw": float(np.cos(_ANGLE)), "sinw": float(np.sin(_ANGLE))},
    "auscom_noslope": {"tilt_mode": 0, "cosw": float(np.cos(_ANGLE)), "sinw": float(np.sin(_ANGLE))},
    "revised": {"revised_evp": True},
}
SEEDS = (1, 2, 3)
STRESSES = tuple(f"stress{k}_{c}" for k in ("p", "m", "12") for c in (1, 2, 3, 4))
IO = ("uvel", "vvel", "fm", "iceumask") + STRESSES
OUT = ("strairx", "strairy", "strtltx", "strtlty", "strocnx", "strocny", "strintx", "strinty", "strocnxT",
       "strocnyT", "divu", "shear", "rdg_conv", "rdg_shear", "prs_sig", "uvel_init", "vvel_init", "tmass",
       "icetmask")


@pytest.fixture(scope="module")
def vec():
    return np.load(NPZ)


def block_case(vec, seed):
    nxb, nyb, ndte, _ncat = (int(v) for v in vec["meta"])
    dom = fx.Domain(nxb - 2, nyb - 2, ew_cyclic=True, ns_tripole=False, ns_open=True, grid_tripole=False,
                    name="ref_block")
    get = lambda n: np.ascontiguousarray(vec[f"in__{seed}__{n}"])
    grid = fx.Grid(dom=dom, f={n: get(n) for n in tuple(fx.GRID_FIELDS) + ("tmask", "umask")})
    inp = {n: get(n) for n in fx.IN_FIELDS}
    io = {n: get(n) for n in fx.INOUT_FIELDS}
    io["iceumask"] = get("iceumask")
    return fx.Case(grid=grid, dt=float(vec["dt"]), ndte=ndte, inp=inp, io=io, out=fx.zeroed_outputs(dom))


def test_vectors_cover_every_branch(vec):
    """The random blocks exercise land, ice-free and ice cells, both iceumask transitions and both signs of fm."""
    for seed in SEEDS:
        icellt, icellu = vec[f"default__{seed}__lists"]
        nxb, nyb = int(vec["meta"][0]), int(vec["meta"][1])
        assert 0 < icellu < icellt < nxb * nyb
        before, after = vec[f"in__{seed}__iceumask"], vec[f"default__{seed}__iceumask"]
        assert ((before == 1) & (after == 0)).any() and ((before == 0) & (after == 1)).any()
        fm = vec[f"default__{seed}__fm"][after == 1]
        assert (fm > 0).any() and (fm < 0).any()
        assert (vec[f"in__{seed}__tmask"] == 0).any()


@pytest.mark.parametrize("seed", SEEDS)
@pytest.mark.parametrize("variant", sorted(VARIANTS))
def test_oracle_reproduces_reference_text_bitwise(vec, variant, seed):
    case = block_case(vec, seed)
    st = orc.run_case(case, no_halo=True, **VARIANTS[variant])
    key = f"{variant}__{seed}__"
    assert (st.active_t_cells, st.active_u_cells) == tuple(int(v) for v in vec[key + "lists"])
    bad = [n for n in IO if not np.array_equal(case.io[n], vec[key + n])]
    bad += [n for n in OUT if not np.array_equal(case.out[n], vec[key + n])]
    assert not bad, f"{variant} seed {seed}: fields differing from the reference text: {bad}"
    sig1, sig2 = orc.principal_stress(case.io["stressp_1"], case.io["stressm_1"], case.io["stress12_1"],
                                      case.out["prs_sig"])
    assert np.array_equal(sig1, vec[key + "sig1"]) and np.array_equal(sig2, vec[key + "sig2"])


def test_variants_differ(vec):
    """The variants are not the same computation five times over."""
    for v in ("coupled", "auscom_slope", "auscom_noslope", "revised"):
        assert not np.array_equal(vec[f"{v}__1__uvel"], vec["default__1__uvel"]), v
    assert not np.array_equal(vec["auscom_slope__1__uvel"], vec["auscom_noslope__1__uvel"])


def test_evp_constants_match_reference_text(vec):
    """set_evp_parameters (ice_dyn_shared.F90:190-240) as executed from the reference text."""
    arlx1i, brlx, denom1, revp, ecci = vec["default__1__consts"]
    dt, ndte = float(vec["dt"]), int(vec["meta"][2])
    dte = dt / ndte
    dte2T = dte / (2.0 * 0.36 * dt)            # tdamp2 = 2*eyc*dt (:216-217)
    assert (arlx1i, brlx, revp, ecci) == (dte2T, dt / dte, 0.0, 0.25)
    assert denom1 == 1.0 / (1.0 + arlx1i)
    arlx1i_r, brlx_r, _d, revp_r, _e = vec["revised__1__consts"]
    assert revp_r == 1.0 and brlx_r > 1.0 and 0.0 < arlx1i_r < 1.0


@pytest.mark.parametrize("seed", SEEDS)
@pytest.mark.parametrize("ks,kp,kr", [(1, 1, 1), (1, 0, 0), (1, 0, 1), (1, 1, 0), (0, 1, 1)])
def test_strength_oracle_reproduces_reference_text_bitwise(vec, seed, ks, kp, kr):
    case = block_case(vec, seed)
    get = lambda n: np.ascontiguousarray(vec[f"in__{seed}__{n}"])
    mask = (vec[f"default__{seed}__icetmask"] == 1).astype(np.int32)
    mine = orc.ice_strength(case.dom, get("aice"), get("vice"), get("aice0"), get("aicen"), get("vicen"),
                            cellmask=mask, kstrength=ks, krdg_partic=kp, krdg_redist=kr)
    ref = vec[f"strength__{seed}__{ks}{kp}{kr}"]
    assert (ref != 0).any()
    assert np.array_equal(mine, ref)
