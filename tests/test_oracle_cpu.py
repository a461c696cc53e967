"""The CPU oracle against (a) an independent numpy restatement written from the Fortran, (b) the
invariants the reference guarantees, (c) its own committed golden vectors.  The reference ships no
golden vectors for this path; tests/test_oracle_reference_text.py pins the
oracle's arithmetic to the reference's subroutine texts instead."""
import os

import math

import numpy as np
import pytest

from cice5_b200 import fixtures as fx
from oracle import oracle as orc, evp_numpy
from tests.util import clone_case, compare_cases, phys

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _run(case, **kw):
    c = clone_case(case)
    orc.run_case(c, **kw)
    return c


@pytest.mark.parametrize("grid,scen,kw", [
    ("gx3", "full", {}),
    ("gx3", "caps", dict(cosw=0.9, sinw=float(np.sqrt(1 - 0.81)), tilt_mode=1, strength_masked=1)),
    ("tp", "full", {}), ("tp", "caps", dict(strength_masked=1)), ("tp_strax", "full", {})])
def test_c_oracle_equals_numpy_restatement_bitwise(grid, scen, kw):
    g = fx.synthetic_tripole(62, 40) if grid.startswith("tp") else fx.make_grid(grid)
    case = fx.make_case(g, scen, ndte=15, dt=3600.0)
    if grid == "tp_strax":         # ACCESS build, calc_strair = .false. (ice_dyn_evp.F90:220-229)
        case.inp["strax"] = 0.7 * case.inp["strairyT"] + 0.01
        case.inp["stray"] = -1.3 * case.inp["strairxT"]
    a = _run(case, **kw)
    b = clone_case(case)
    nt, nu = evp_numpy.evp(b, **kw)
    compare_cases(b, a, bitwise=True)
    assert nu == int(phys(a.io["iceumask"]).sum())


@pytest.mark.parametrize("grid,slabs", [("gx3", (2, 4, 5)), ("tp", (2, 4, 8))])
def test_decomposition_independence_bitwise(grid, slabs):
    """cicedoc.tex:4601-4603: results do not depend on the block decomposition."""
    g = fx.synthetic_tripole(96, 40) if grid == "tp" else fx.make_grid(grid)
    case = fx.make_case(g, "full", ndte=30, dt=3600.0)
    ref = _run(case, nslabs=1)
    for n in slabs:
        compare_cases(_run(case, nslabs=n), ref, bitwise=True)


def test_tripole_seam_antisymmetry():
    """After evp the top U row is antisymmetric under i -> nx - i."""
    g = fx.synthetic_tripole(96, 40)
    c = _run(fx.make_case(g, "full", ndte=60, dt=3600.0))
    nx, ny = g.dom.nx, g.dom.ny
    for f in ("uvel", "vvel"):
        top = c.io[f][ny, 1:nx + 1]                    # ig = 1..nx
        i = np.arange(1, nx // 2)
        assert np.array_equal(top[i - 1], -top[nx - i - 1])
        assert np.abs(top).max() > 0.0
        # ghost row = minus the mirrored row ny-1
        ghost, below = c.io[f][ny + 1, 1:nx + 1], c.io[f][ny - 1, 1:nx + 1]
        ig = np.arange(1, nx)
        assert np.array_equal(ghost[ig - 1], -below[nx - ig - 1])


def test_zero_forcing_zero_state_stays_at_rest():
    case = fx.make_case("gx3", "full", ndte=20)
    for k in ("strairxT", "strairyT", "uocn", "vocn"):
        case.inp[k][...] = 0.0
    c = _run(case)
    for k in ("uvel", "vvel") + fx.STRESS_FIELDS:
        assert np.abs(phys(c.io[k])).max() == 0.0


def test_free_drift_matches_closed_form_solve():
    """strength = 0 => no internal stress; each subcycle is the pointwise 2x2 solve of stepu
    (ice_dyn_shared.F90:708-737) -- iterated here in numpy from the oracle's own prep outputs."""
    case = fx.make_case("gx3", "full", ndte=25)
    case.inp["strength"][...] = 0.0
    c = _run(case)
    ice = c.io["iceumask"] > 0
    P = (slice(1, -1), slice(1, -1))
    g, inp = case.grid.f, case.inp
    tmass = c.out["tmass"]
    umass = evp_numpy.to_ugrid(tmass, g["tarea"], g["uarea"])
    aiu = evp_numpy.to_ugrid(inp["aice_init"], g["tarea"], g["uarea"])
    fm = c.io["fm"]
    forcex = c.out["strairx"] + c.out["strtltx"]
    forcey = c.out["strairy"] + c.out["strtlty"]
    u, v = c.out["uvel_init"].copy(), c.out["vvel_init"].copy()
    brlx = evp_numpy.evp_parameters(case.dt, case.ndte)[2]
    for _ in range(case.ndte):
        vrel = aiu * fx.RHOW * inp["Cw"] * np.sqrt((inp["uocn"] - u) * (inp["uocn"] - u) + (inp["vocn"] - v) * (inp["vocn"] - v))
        cca = brlx * (umass / case.dt) + vrel * 1.0
        ccb = fm + np.copysign(1.0, fm) * vrel * 0.0
        cc1 = 0.0 + forcex + vrel * inp["uocn"] + (umass / case.dt) * (brlx * u)
        cc2 = 0.0 + forcey + vrel * inp["vocn"] + (umass / case.dt) * (brlx * v)
        ab2 = cca * cca + ccb * ccb
        with np.errstate(divide="ignore", invalid="ignore"):
            un, vn = (cca * cc1 + ccb * cc2) / ab2, (cca * cc2 - ccb * cc1) / ab2
        u, v = np.where(ice, un, u), np.where(ice, vn, v)
    assert np.allclose(u[P], c.io["uvel"][P], rtol=1e-12, atol=1e-15)
    assert np.allclose(v[P], c.io["vvel"][P], rtol=1e-12, atol=1e-15)
    assert np.abs(phys(c.io["stressp_1"])).max() == 0.0


@pytest.mark.parametrize("tripole", [True, False])
@pytest.mark.parametrize("loc,kind", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_halo_update_c_vs_numpy(tripole, loc, kind):
    """ice_HaloUpdate (center / NEcorner x scalar / vector), C oracle with 1..4 blocks against the
    numpy implementation in fixtures.halo_update."""
    g = fx.synthetic_tripole(48, 20) if tripole else fx.synthetic_displaced(48, 20)
    rng = np.random.default_rng(3)
    a = rng.standard_normal((22, 50))
    ref = fx.halo_update(g.dom, a.copy(), loc, kind, 0.0)
    for n in (1, 2, 3, 4):
        got = orc.halo_r8(g.dom, a.copy(), loc, kind, 0.0, nslabs=n)
        assert np.array_equal(got, ref), n


def test_principal_stress_diagnostic():
    rng = np.random.default_rng(0)
    sp, sm, s12 = (rng.standard_normal(100) for _ in range(3))
    prs = np.abs(rng.standard_normal(100)); prs[:5] = 0.0
    s1, s2 = orc.principal_stress(sp, sm, s12, prs)
    assert np.all(s1[:5] == 1e30) and np.all(s2[:5] == 1e30)
    r = np.sqrt(sm * sm + 4.0 * (s12 * s12))
    assert np.array_equal(s1[5:], ((0.5 * (sp + r)) / prs)[5:])


@pytest.mark.parametrize("name,grid", [("gx3_full", "gx3"), ("tp62x40_full", None)])
def test_oracle_reproduces_committed_golden_vectors(name, grid):
    z = np.load(os.path.join(GOLD, f"{name}_ndte120_oracle.npz"))
    g = fx.make_grid(grid) if grid else fx.synthetic_tripole(62, 40)
    c = _run(fx.make_case(g, "full", ndte=120, dt=3600.0), nslabs=2)
    for k in z.files:
        grp, f = k.split("_", 1)
        assert np.array_equal(getattr(c, grp if grp == "io" else "out")[f], z[k]), k


def test_umax_sanity():
    """check_umax (ice_diagnostics.F90:340-364): |u| stays below umax_stab = 1 m/s."""
    c = _run(fx.make_case("gx3", "full", ndte=120))
    assert np.hypot(c.io["uvel"], c.io["vvel"]).max() < 1.0


# ---- revised EVP (ice_dyn_shared.F90:225-233,492-504,715,732-734) ----------

@pytest.mark.parametrize("grid", ["gx3", "tp"])
def test_revised_evp_c_oracle_equals_numpy_and_is_slab_independent(grid):
    g = fx.synthetic_tripole(62, 40) if grid == "tp" else fx.make_grid(grid)
    case = fx.make_case(g, "full", ndte=15, dt=3600.0)
    a = _run(case, revised_evp=True)
    b = clone_case(case)
    evp_numpy.evp(b, revised_evp=True)
    compare_cases(b, a, bitwise=True)
    compare_cases(_run(case, revised_evp=True, nslabs=2), a, bitwise=True)
    classic = _run(case)
    assert not np.array_equal(phys(a.io["uvel"]), phys(classic.io["uvel"]))      # really another scheme


def test_revised_evp_parameters_and_stress_reset():
    """revp = 1, 1/alpha = 2 xi / Se, beta = 2 Se xi gamma / xmin^2 with xmin over tmask cells; the 12
    stresses restart from zero on every call (evp_prep2 :492-504), so a second call with the same
    forcing does not depend on the stresses the first one left."""
    g = fx.make_grid("gx3")
    tm = g.f["tmask"][1:-1, 1:-1] > 0
    xmin = min(g.f["dxt"][1:-1, 1:-1][tm].min(), g.f["dyt"][1:-1, 1:-1][tm].min())
    ecci, arlx1i, brlx, denom1 = evp_numpy.evp_parameters(3600.0, 120, True, xmin)
    assert arlx1i == 2.0 * 5.5e-3 / 0.86 and denom1 == 1.0 / (1.0 + arlx1i)
    assert brlx == 2.0 * 0.86 * 5.5e-3 * (0.25 * 1.0e11 * 3600.0) / (xmin * xmin)
    case = fx.make_case(g, "full", ndte=20)
    a = _run(case, revised_evp=True)
    b = clone_case(a)
    for n in fx.STRESS_FIELDS:
        b.io[n] = np.random.default_rng(1).normal(size=b.io[n].shape) * 1.0e3     # garbage stresses in
    b.out = fx.zeroed_outputs(case.dom)
    a.out = fx.zeroed_outputs(case.dom)
    orc.run_case(a, revised_evp=True)
    orc.run_case(b, revised_evp=True)
    compare_cases(b, a, bitwise=True)


# ---- ice_strength (ice_mechred.F90:2111-2269, ridge_itd :936-1285) ---------

def _itd_case(grid, scen):
    g = fx.make_grid(grid)
    st = fx.ice_state(g, scen)
    return g, st, fx.ice_itd(st)


@pytest.mark.parametrize("grid,scen", [("gx3", "full"), ("tp1", "caps")])
@pytest.mark.parametrize("partic,redist", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_strength_c_oracle_equals_numpy_restatement(grid, scen, partic, redist):
    from oracle import strength_numpy as sn
    g, st, itd = _itd_case(grid, scen)
    a = orc.ice_strength(g.dom, st["aice"], st["vice"], itd["aice0"], itd["aicen"], itd["vicen"],
                         krdg_partic=partic, krdg_redist=redist)
    b = sn.ice_strength(st["aice"], st["vice"], itd["aice0"], itd["aicen"], itd["vicen"],
                        krdg_partic=partic, krdg_redist=redist, rhow=fx.RHOW, rhoi=fx.RHOI, gravit=fx.GRAVIT)
    assert np.isfinite(a).all() and (a >= 0).all() and a.max() > 1.0e3
    if partic == 0:
        assert np.array_equal(a, b)                       # no exp: the same IEEE operations
    else:                                                 # numpy's vector exp vs libm: last-bit differences
        assert np.max(np.abs(a - b) / np.maximum(np.abs(a), 1e-300)) < 4e-16


def test_cr_exp_is_correctly_rounded():
    """evp_exp (csrc/evp_exp.h: the one exp the device kernels and the oracle share) against exact rational
    arithmetic: decimal's exp at 70 digits, rounded to the nearest double by Fraction.__float__.  Arguments:
    the path's domain [-20, 0] densely, the whole normal range sparsely, the structured values ridge_itd
    produces (-G/astar with G = k/n), and edge values."""
    from decimal import Decimal, getcontext
    from fractions import Fraction
    getcontext().prec = 70
    rng = np.random.default_rng(20261105)
    xs = np.concatenate([rng.uniform(-20.0, 0.0, 6000), rng.uniform(-700.0, 700.0, 1500),
                         -np.arange(0, 1001, 7) / 1000.0 / 0.05, -rng.uniform(0, 1, 500) * (1.0 / 0.05),
                         np.array([0.0, -0.0, 1.0, -1.0, -20.0, 1e-300, -1e-300, 1e-17, -1e-17, 0.5, -0.5,
                                   2.0 ** -30, -(2.0 ** -30), 88.0, -88.0, 699.5, -699.5])])
    got = orc.cr_exp(xs)
    want = np.array([float(Fraction(Decimal(float(x)).exp())) for x in xs])
    assert np.array_equal(got, want), f"{np.count_nonzero(got != want)} of {xs.size} not correctly rounded"
    # libm's exp is within one ulp of it (and not always equal: that is why the device cannot use "the" exp)
    lm = np.array([math.exp(float(x)) for x in xs])
    assert np.max(np.abs(lm - want) / np.spacing(want)) <= 1.0


@pytest.mark.parametrize("partic,redist,ks", [(1, 1, 1), (1, 0, 1), (1, 1, 0)])
def test_strength_with_cr_exp_within_ulps_of_libm_build(partic, redist, ks):
    """exp_mode 1 (evp_exp, the device's) against exp_mode 0 (libm, what the reference-text vectors pin):
    the same statements, exp differing by at most one ulp in about one argument in a thousand."""
    g, st, itd = _itd_case("tp1", "caps")
    kw = dict(kstrength=ks, krdg_partic=partic, krdg_redist=redist)
    a = orc.ice_strength(g.dom, st["aice"], st["vice"], itd["aice0"], itd["aicen"], itd["vicen"], **kw)
    b = orc.ice_strength(g.dom, st["aice"], st["vice"], itd["aice0"], itd["aicen"], itd["vicen"], cr_exp=True, **kw)
    assert np.array_equal(a == 0, b == 0) and b.max() > 1.0e3
    assert np.max(np.abs(a - b) / np.maximum(np.abs(a), 1e-300)) < 2e-14


def test_strength_list_semantics_hibler_and_itd_fixture():
    g, st, itd = _itd_case("gx3", "caps")
    assert np.allclose(itd["aicen"].sum(axis=0), st["aice"], rtol=0, atol=1e-15)
    assert np.allclose(itd["vicen"].sum(axis=0), st["vice"], rtol=1e-14, atol=0)
    mask = (st["aice"] > 0.5).astype(np.int32)
    full = orc.ice_strength(g.dom, st["aice"], st["vice"], itd["aice0"], itd["aicen"], itd["vicen"])
    lst = orc.ice_strength(g.dom, st["aice"], st["vice"], itd["aice0"], itd["aicen"], itd["vicen"],
                           cellmask=mask)
    assert np.array_equal(lst, np.where(mask != 0, full, 0.0))      # 0 outside the list (:2175)
    hib = orc.ice_strength(g.dom, st["aice"], st["vice"], itd["aice0"], itd["aicen"], itd["vicen"],
                           kstrength=0)
    # fixtures' Hibler strength (:2263); numpy's vector exp and libm's differ in the last bit
    assert np.allclose(phys(hib), phys(st["strength"]), rtol=4e-16, atol=0) and hib.max() > 1e4
    assert not hib[0, :].any() and not hib[:, 0].any()               # physical cells only


def test_strength_single_category_closed_form():
    """One category of thickness h under 100 % cover, Thorndike participation, Hibler 80
    redistribution: G = (0, 0, 1), apartic(1) = 1, so
    strength = Cf Cp (-h^2 + (hmax^3 - hmin^3) / (3 (hmax - hmin)) / k) / (1 - 1/k)."""
    g = fx.make_grid("gx3")
    shape = g.f["tmask"].shape
    h = 2.0
    aicen = np.ones((1,) + shape)
    vicen = np.full((1,) + shape, h)
    s = orc.ice_strength(g.dom, np.ones(shape), vicen[0], np.zeros(shape), aicen, vicen,
                         krdg_partic=0, krdg_redist=0)
    hmin = min(2 * h, h + 1.0)
    hmax = max(2 * np.sqrt(25.0 * h), hmin + 1e-11)
    k = 0.5 * (hmin + hmax) / h
    Cp = 0.5 * fx.GRAVIT * (fx.RHOW - fx.RHOI) * fx.RHOI / fx.RHOW
    want = 17.0 * Cp * (-h * h + (hmax ** 3 - hmin ** 3) / (3 * (hmax - hmin)) / k) / (1 - 1 / k)
    assert abs(s[5, 5] - want) <= 1e-12 * want


def test_wet_pole_tripole_velocity_fold_skips_the_pole_pair():
    """The velocity pair goes through the reference's 3-D halo routine, whose NE-corner fold does not
    average the column pair (1, nx-1) next to the pole (oracle/evp_oracle.c halo_r8_ex; pinned to the
    reference text by tests/test_oracle_reference_text.py).  On real grids those U points are land; the
    `wet_poles` fixture puts ice on them: C oracle == numpy restatement bit for bit, 1 == 2 == 4 slabs,
    every other pair of the top row antisymmetric, the pole pair swapped with its sign but not averaged."""
    g = fx.synthetic_tripole(48, 24, wet_poles=True)
    nx, ny = g.dom.nx, g.dom.ny
    assert g.f["umask"][ny, 1] and g.f["umask"][ny, nx - 1]
    case = fx.make_case(g, "full", ndte=12, dt=3600.0)
    a = _run(case)
    b = clone_case(case)
    evp_numpy.evp(b)
    compare_cases(b, a, bitwise=True)
    for n in (2, 4):
        compare_cases(_run(case, nslabs=n), a, bitwise=True)
    for f in ("uvel", "vvel"):
        top = a.io[f][ny, 1:nx + 1]
        i = np.arange(2, nx // 2)
        assert np.array_equal(top[i - 1], -top[nx - i - 1])
        assert top[0] != 0.0 and top[0] != -top[nx - 2]


def test_shared_trig_accuracy_against_multiprecision():
    """evp_trig.h (the sin / cos / atan2 the EAP kernels and the oracle share) against 200-bit arithmetic:
    sin, cos < 1 ulp for |x| <= 1e5, atan2 < 1.5 ulp; C's conventions for signed zeros and the axes."""
    import math
    mp = pytest.importorskip("mpmath")
    mp.mp.prec = 200
    fsin, fcos, fatan2 = orc.trig("sin"), orc.trig("cos"), orc.trig("atan2")

    def ulps(got, exact):
        return float(abs(mp.mpf(got) - exact) / math.ulp(float(exact)))

    r = np.random.default_rng(7)
    xs = np.concatenate([r.uniform(-4, 4, 4000), r.uniform(-1e5, 1e5, 1500), r.uniform(-1e-3, 1e-3, 500),
                         [k * math.pi / 4 for k in range(-16, 17)]])
    worst_s = max(ulps(fsin(float(x)), mp.sin(float(x))) for x in xs)
    worst_c = max(ulps(fcos(float(x)), mp.cos(float(x))) for x in xs)
    assert worst_s < 1.0 and worst_c < 1.0, (worst_s, worst_c)
    worst_a = 0.0
    for _ in range(6000):
        y, x = float(r.normal() * 10 ** r.uniform(-6, 6)), float(r.normal() * 10 ** r.uniform(-6, 6))
        worst_a = max(worst_a, ulps(fatan2(y, x), mp.atan2(y, x)))
    for _ in range(6000):
        y, x = float(r.uniform(-1, 1)), float(r.uniform(-1, 1))
        worst_a = max(worst_a, ulps(fatan2(y, x), mp.atan2(y, x)))
    assert worst_a < 1.5, worst_a
    for y, x in ((0.0, 1.0), (-0.0, 1.0), (0.0, -1.0), (-0.0, -1.0), (0.0, 0.0), (-0.0, 0.0), (0.0, -0.0), (-0.0, -0.0),
                 (1.0, 0.0), (-1.0, 0.0), (1.0, -0.0), (1.0, 1.0), (1.0, -1.0), (-1.0, -1.0), (-1.0, 1.0)):
        g, e = fatan2(y, x), math.atan2(y, x)
        assert g == e and math.copysign(1.0, g) == math.copysign(1.0, e), (y, x, g, e)
    assert fsin(0.0) == 0.0 and fcos(0.0) == 1.0 and math.copysign(1.0, fsin(-0.0)) == -1.0
    # the one constant of update_stress_rdg that goes through sin: sin(pi2/c12) (ice_dyn_eap.F90:1525)
    assert fsin(2.0 * 3.14159265358979323846 / 12.0) == math.sin(2.0 * 3.14159265358979323846 / 12.0)


@pytest.mark.parametrize("slabs", [2, 4])
def test_eap_decomposition_independence_and_invariants(slabs):
    """eap(dt) of the oracle: x-slabs give the bits of one block; the structure tensor stays a valid one
    (eigenvalues of [[a11, a12], [a12, 1 - a11]] in [0, 1]) and isotropic where there is no ice."""
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    case = fx.make_case(grid, "full", ndte=21, dt=3600.0)
    tables = orc.eap_tables()
    res = []
    for ns in (1, slabs):
        c = clone_case(case)
        e = orc.eap_state(case.dom)
        orc.run_case_eap(c, e, tables, nslabs=ns)
        res.append((c, e))
    (c1, e1), (c2, e2) = res
    for n in ("uvel", "vvel", "stressp_1", "stress12_3"):
        assert np.array_equal(phys(c1.io[n]), phys(c2.io[n])), n
    for n in orc.EAP_CORNER + orc.EAP_AVG + orc.EAP_OUT:
        assert np.array_equal(phys(e1[n]), phys(e2[n])), n
    ice = phys(c1.out["icetmask"]) == 1
    assert ice.any() and (~ice).any()
    for c in (1, 2, 3, 4):
        a11, a12 = phys(e1[f"a11_{c}"]), phys(e1[f"a12_{c}"])
        lam = 0.5 + np.sqrt((a11 - 0.5) ** 2 + a12 ** 2)
        assert (lam <= 1.0 + 1e-12).all() and (a11[~ice] == 0.5).all() and (a12[~ice] == 0.0).all()
        assert (a11[ice] != 0.5).any()            # stepa moved it (subcycles 1, 11, 21)
    assert np.abs(phys(c1.io["uvel"])).max() < 5.0


def test_fixture_eap_tables_agree_with_the_oracles_generator():
    """fixtures.eap_tables (vectorised numpy, what bench.py and tools/ hand to evp_gpu_set_eap_tables) against the oracle's
    restatement of init_eap's loop (pinned bit for bit to the reference text on 64 entries): equal to the rounding of the
    ill-conditioned weight polynomials w1, w2 (coefficients of 1e6 cancelling to order 1)."""
    T, R = fx.eap_tables(), orc.eap_tables()
    for n in fx.EAP_TABLES:
        assert T[n].shape == orc.EAP_TABLE_SHAPE
        d = np.abs(T[n] - R[n])
        flipped = (T[n] == 0.0) != (R[n] == 0.0)                  # |value| within rounding of the 1e-6 cut-off
        assert d[~flipped].max() < 1.0e-9 and flipped.sum() <= 4 and d.max() <= 1.0e-6 + 1.0e-9, (n, d.max(), flipped.sum())
        assert np.abs(R[n]).max() > 0.1
