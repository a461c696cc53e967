"""The EAP oracle against vectors produced by the REFERENCE'S OWN TEXT (source/ice_dyn_eap.F90).

tests/golden/ref_eap_vectors.npz was written by tests/golden/make_eap_vectors.py (fortran_exec.py, the
Fortran-subset interpreter): w1, w2, the six table kernels, init_eap's table loop on 64 entries, update_stress_rdg,
calc_ffrac, and the driver's call order on seeded 12 x 10 blocks (evp_prep1 ... ndte x (stress_eap, stepu, stepa) ...
evp_finish, to_tgrid; halo updates left out).  The C oracle has to reproduce every value BIT FOR BIT.  The intrinsics
atan2 / sin / cos of the in-loop routines are cice5_b200/csrc/evp_trig.h on both sides (the generator's docstring
says why); the table code runs with the C library's functions on both sides.  Nothing here reads /root/reference.
"""
import ctypes as C
import math
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests.test_oracle_reference_text import IO, OUT, block_case

NPZ = os.path.join(os.path.dirname(__file__), "golden", "ref_eap_vectors.npz")
BLOCK_NPZ = os.path.join(os.path.dirname(__file__), "golden", "ref_block_vectors.npz")
SEEDS = (1, 2)
EAP_FIELDS = orc.EAP_CORNER + orc.EAP_AVG + orc.EAP_OUT


@pytest.fixture(scope="module")
def vec():
    return np.load(NPZ)


@pytest.fixture(scope="module")
def block_vec():
    return np.load(BLOCK_NPZ)


def test_w_and_table_kernels_bitwise(vec):
    lib = orc.lib()
    lib.eap_oracle_w.restype = C.c_double
    lib.eap_oracle_table_kernel.restype = C.c_double
    a = vec["unit__w_a"]
    for which in (1, 2):
        mine = np.array([lib.eap_oracle_w(which, C.c_double(v)) for v in a])
        assert np.array_equal(mine, vec[f"unit__w{which}"])
    phi = math.pi / 12.0
    for k in range(6):
        mine = np.array([lib.eap_oracle_table_kernel(k, C.c_double(x), C.c_double(y), C.c_double(z), C.c_double(phi))
                         for x, y, z in vec["unit__kernel_xyz"]])
        ref = vec[f"unit__kernel_{k}"]
        assert np.array_equal(mine, ref) and np.array_equal(np.signbit(mine), np.signbit(ref)), k
        assert len(np.unique(ref)) > 10                      # not a degenerate sample


def test_lookup_table_entries_bitwise(vec):
    """init_eap's table loop (source/ice_dyn_eap.F90:567-618) on 64 entries incl. the corners and ia = na_yield."""
    T = orc.eap_tables()
    entries = vec["unit__table_entries"]
    assert len(entries) == 64 and any(e[0] == 21 for e in entries)
    nonzero = 0
    for n in orc.EAP_TABLES:
        mine = np.array([T[n][ia - 1, iy - 1, ix - 1] for ia, ix, iy in entries])
        assert np.array_equal(mine, vec[f"unit__table_{n}"]), n
        nonzero += int(np.count_nonzero(mine))
    assert nonzero > 100
    # a partial fill writes only its range
    part = orc.eap_tables(ia=(3, 3), ix=(5, 6), iy=(7, 7))
    assert np.count_nonzero(part["s11r"]) <= 2 and part["s11r"][2, 6, 4] == T["s11r"][2, 6, 4]


def test_update_stress_rdg_bitwise(vec):
    lib = orc.lib()
    T = orc.eap_tables()
    E = orc._eap_struct(T, orc.eap_state(type("D", (), {"nx": 2, "ny": 2})()), trig_mode=1)
    f = lib.eap_oracle_update_stress_rdg
    f.restype = None
    arg = {k: vec[f"unit__rdg_{k}"] for k in ("divu", "tension", "shear", "a11", "a12", "strength")}
    ref = vec["unit__rdg_out"]
    kx_seen = set()
    for i in range(len(ref)):
        for last in (0, 1):
            o = [C.c_double(-7.0) for _ in range(5)]
            f(C.byref(E), 12 if last else 3, 12, *[C.c_double(arg[k][i]) for k in
                                                   ("divu", "tension", "shear", "a11", "a12", "strength")],
              *[C.byref(v) for v in o])
            mine = np.array([v.value for v in o])
            assert np.array_equal(mine, ref[i, last]), (i, last, mine, ref[i, last])
        kx_seen.add(round(float(ref[i, 0, 0]), 3))
    assert len(kx_seen) > 100                                # many different table entries were hit
    assert (ref[:, 0, 3:] == -7.0).all() and (ref[:, 1, 3:] != -7.0).all()   # alphar / alphas only when ksub == ndte


def test_calc_ffrac_bitwise(vec):
    f = orc.lib().eap_oracle_calc_ffrac
    f.restype = C.c_double
    arg = {k: vec[f"unit__ffrac_{k}"] for k in ("stressp", "stressm", "stress12", "a1x")}
    ref = vec["unit__ffrac_out"]
    for i in range(len(ref)):
        for b in (1, 2):
            mine = f(1, b, *[C.c_double(arg[k][i]) for k in ("stressp", "stressm", "stress12", "a1x")])
            assert mine == ref[i, b - 1], (i, b)
    assert (ref == 0.0).any() and (ref != 0.0).any()         # both kinds of branch


@pytest.mark.parametrize("seed", SEEDS)
def test_eap_oracle_reproduces_reference_text_bitwise(vec, block_vec, seed):
    case = block_case(block_vec, seed)
    case.ndte = int(vec["meta"][2])
    eap = {n: np.ascontiguousarray(vec[f"in__{seed}__{n}"]) for n in EAP_FIELDS}
    st = orc.run_case_eap(case, eap, no_halo=True, trig_mode=1)
    key = f"block__{seed}__"
    assert (st.active_t_cells, st.active_u_cells) == tuple(int(v) for v in vec[key + "lists"])
    bad = [n for n in IO if not np.array_equal(case.io[n], vec[key + n])]
    bad += [n for n in OUT if not np.array_equal(case.out[n], vec[key + n])]
    bad += [n for n in EAP_FIELDS if not np.array_equal(eap[n], vec[key + n])]
    assert not bad, f"seed {seed}: fields differing from the reference text: {bad}"
    for n in ("rdg_conv", "yieldstress12", "a12_2"):         # signs of zeros too (rdg_conv = -min(x, 0) * tarear)
        mine = case.out[n] if n in case.out else eap[n]
        assert np.array_equal(np.signbit(mine), np.signbit(vec[key + n])), n
    # stepa ran (subcycles 1 and 11) and the structure tensor was reset off the ice
    assert not np.array_equal(vec[key + "a11_1"], vec[f"in__{seed}__a11_1"])
    off = vec[key + "icetmask"] == 0
    assert off.any() and (vec[key + "a11_3"][off] == 0.5).all() and (vec[key + "a12_3"][off] == 0.0).all()
    assert np.count_nonzero(vec[key + "yieldstress11"]) > 10 and np.signbit(vec[key + "rdg_conv"]).any()
    assert not vec[key + "rdg_shear"].any()                  # commented out in the reference (:1283-1285)


def test_libm_intrinsics_change_little(vec, block_vec):
    """trig_mode = 0 (the C library's atan2 / sin / cos) against trig_mode = 1: the same physics; the two may differ
    where a last bit moves a cell to the neighbouring table entry, which is why the device shares evp_trig.h."""
    res = []
    for mode in (0, 1):
        case = block_case(block_vec, 1)
        case.ndte = int(vec["meta"][2])
        eap = {n: np.ascontiguousarray(vec[f"in__1__{n}"]) for n in EAP_FIELDS}
        orc.run_case_eap(case, eap, no_halo=True, trig_mode=mode)
        res.append((case.io["uvel"].copy(), eap["a11_1"].copy()))
    scale = np.abs(res[1][0]).max()
    assert np.abs(res[0][0] - res[1][0]).max() <= 1e-6 * scale
    assert np.abs(res[0][1] - res[1][1]).max() <= 1e-9
