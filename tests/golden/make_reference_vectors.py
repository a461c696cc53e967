#!/usr/bin/env python
"""Golden vectors from the REFERENCE'S OWN SUBROUTINE TEXTS (run in the build container only).

The reference cannot be compiled here (no Fortran compiler), so its arithmetic subroutines are
executed mechanically by tests/golden/fortran_exec.py straight from the checkout at
/root/reference:

  source/ice_dyn_shared.F90   set_evp_parameters, evp_prep1, evp_prep2, stepu, evp_finish,
                              principal_stress
  source/ice_dyn_evp.F90      stress
  source/ice_grid.F90         to_ugrid, to_tgrid
  source/ice_mechred.F90      ice_strength, asum_ridging, ridge_itd
  drivers/cice/ice_constants.F90  every `parameter`

on one small block with seeded random inputs, in the order the driver `evp` calls them
(source/ice_dyn_evp.F90:150-506) but WITHOUT the halo updates (ghost cells stay as given).  Inputs
and outputs go to tests/golden/ref_block_vectors.npz; tests/test_oracle_cpu.py then demands that
the C oracle in its `no_halo` mode reproduces every output bit for bit.  This pins the oracle's
arithmetic (a2-a7, a10-a12 and the strength of SURVEY.md section 8a) to the reference text; what
stays restated by hand is the driver's call order, the halo / tripole code and the list building
order of the C side (the lists themselves come out of the reference's evp_prep2 here).

    python tests/golden/make_reference_vectors.py        # rewrites the .npz
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import fortran_exec as fe  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(HERE, "ref_block_vectors.npz")

NXB, NYB = 12, 10            # block incl. the ghost ring
NDTE, DT = 4, 3600.0
NCAT = 5


def load_modules(defined=(), overrides=None):
    """Parse the constants and translate the subroutines; `defined` = CPP macros that are set."""
    def lines(rel):
        return fe.preprocess(open(os.path.join(REF, rel)).read(), defined)

    ns = {}
    fe.parse_parameters(lines("drivers/cice/ice_constants.F90"), ns)
    shared, evp_l, grid_l, mech = (lines("source/ice_dyn_shared.F90"), lines("source/ice_dyn_evp.F90"),
                                   lines("source/ice_grid.F90"), lines("source/ice_mechred.F90"))
    fe.parse_parameters(shared, ns)          # a_min, m_min, eyc, cosw, sinw (non-AusCOM)
    fe.parse_parameters(mech, ns)            # Cp, Gstar, astar, maxraft, Hstar, Pstar, Cstar
    if overrides:
        ns.update(overrides)
    m = fe.Module(ns)
    m.load(shared, ["set_evp_parameters", "evp_prep1", "evp_prep2", "stepu", "evp_finish", "principal_stress"])
    m.load(evp_l, ["stress"])
    m.load(grid_l, ["to_ugrid", "to_tgrid"])
    m.load(mech, ["asum_ridging", "ridge_itd", "ice_strength"])
    return m


def random_block(seed):
    """Inputs of one block; (NYB, NXB) arrays, C order == Fortran (nx_block, ny_block)."""
    r = np.random.default_rng(seed)
    sh = (NYB, NXB)
    u = lambda lo, hi: r.uniform(lo, hi, sh)
    g = {}
    g["dxt"], g["dyt"] = u(2.0e4, 3.0e4), u(2.0e4, 3.0e4)
    g["dxhy"], g["dyhx"] = u(-8.0e2, 8.0e2), u(-8.0e2, 8.0e2)
    g["cxp"], g["cyp"] = u(2.0e4, 3.0e4), u(2.0e4, 3.0e4)
    g["cxm"], g["cym"] = -u(2.0e4, 3.0e4), -u(2.0e4, 3.0e4)
    g["tarea"] = g["dxt"] * g["dyt"]
    g["uarea"] = u(4.0e8, 9.0e8)
    g["tarear"], g["uarear"] = 1.0 / g["tarea"], 1.0 / g["uarea"]
    g["tinyarea"] = 1.0e-11 * g["tarea"]
    g["fcor"] = u(-1.4e-4, 1.4e-4)
    tm = r.uniform(size=sh) > 0.12
    um = np.zeros(sh, bool)
    um[:-1, :-1] = tm[:-1, :-1] & tm[:-1, 1:] & tm[1:, :-1] & tm[1:, 1:]
    g["tmask"], g["umask"] = tm.astype(np.int32), um.astype(np.int32)
    s = {}
    s["aice"] = np.where(r.uniform(size=sh) > 0.2, u(0.05, 1.0), u(0.0, 2.0e-3))
    s["vice"] = s["aice"] * u(0.3, 3.0)
    s["vsno"] = s["aice"] * u(0.0, 0.4)
    s["aice_init"] = np.clip(s["aice"] + u(-0.02, 0.02), 0.0, 1.0)
    s["strairxT"], s["strairyT"] = u(-0.3, 0.3), u(-0.3, 0.3)
    s["uocn"], s["vocn"] = u(-0.2, 0.2), u(-0.2, 0.2)
    s["ss_tltx"], s["ss_tlty"] = u(-1.0e-5, 1.0e-5), u(-1.0e-5, 1.0e-5)
    s["Cw"] = u(4.0e-3, 7.0e-3)
    s["strength"] = u(1.0e3, 5.0e4) * (s["aice"] > 0.01)
    s["uvel"], s["vvel"] = u(-0.3, 0.3), u(-0.3, 0.3)
    for k in ("p", "m", "12"):
        for c in (1, 2, 3, 4):
            s[f"stress{k}_{c}"] = u(-2.0e3, 2.0e3)
    s["iceumask"] = (r.uniform(size=sh) > 0.5).astype(np.int32)
    s["fm"] = u(-1.0, 1.0)                       # overwritten on ice points, kept elsewhere
    # thickness distribution for ice_strength
    w = np.array([0.1, 0.2, 0.3, 0.25, 0.15])
    jitter = r.uniform(0.6, 1.4, (NCAT,) + sh)
    aicen = w[:, None, None] * jitter * s["aice"]
    aicen *= np.minimum(1.0, 0.999 / np.maximum(aicen.sum(axis=0), 1e-30))
    hn = np.array([0.3, 0.7, 1.0, 1.5, 2.5])[:, None, None] * r.uniform(0.8, 1.2, (NCAT,) + sh)
    s["aicen"], s["vicen"] = aicen, aicen * hn
    s["aice0"] = 1.0 - aicen.sum(axis=0)
    return g, s


def run_reference_sequence(m, g, s, revised_evp=False, ndte=None):
    """The driver's call order (source/ice_dyn_evp.F90:150-506) on one block, halos left out.
    Returns {name: (NYB, NXB) array} of every inout / out field."""
    NDTE = ndte or globals()["NDTE"]
    A = fe.FArr.from_numpy
    ns = m.ns
    ilo, ihi, jlo, jhi = 2, NXB - 1, 2, NYB - 1
    blk = types.SimpleNamespace(ilo=ilo, ihi=ihi, jlo=jlo, jhi=jhi)
    F = {k: A(v) for k, v in g.items() if k not in ("tmask", "umask")}
    tmask, umask = A(g["tmask"] > 0), A(g["umask"] > 0)
    # module-level names the subroutines use
    ns.update(ndte=NDTE, revised_evp=revised_evp, my_task=0, master_task=0, distrb_info=None,
              dxt=F["dxt"], dyt=F["dyt"], tmask=tmask, nblocks=1, blocks_ice=lambda i: i,
              get_block=lambda *_: blk,
              global_minval=lambda arr, _d, msk: float(arr.a[1:-1, 1:-1][msk.a[1:-1, 1:-1]].min()))
    m.call("set_evp_parameters", dt=DT)          # sets the module variables dtei, ecci, dte2t, revp, arlx1i, ...
    k = ns

    z = lambda: fe.FArr([(1, NXB), (1, NYB)])
    zi = lambda: fe.FArr([(1, NXB), (1, NYB)], dtype=int)
    S = {n: A(s[n]) for n in ("aice", "vice", "vsno", "aice_init", "strairxT", "strairyT", "uocn", "vocn",
                              "ss_tltx", "ss_tlty", "Cw", "strength", "uvel", "vvel", "fm")}
    ST = {n: A(s[n]) for n in s if n.startswith("stress")}
    iceumask = A(s["iceumask"] > 0)
    O = {n: z() for n in ("strairx", "strairy", "strtltx", "strtlty", "strocnx", "strocny", "strintx", "strinty",
                          "strocnxT", "strocnyT", "divu", "shear", "rdg_conv", "rdg_shear", "prs_sig",
                          "uvel_init", "vvel_init", "tmass", "umass", "aiu", "umassdti", "waterx", "watery",
                          "forcex", "forcey")}
    icetmask = zi()
    # evp :153-170 (the zeroing loop), then evp_prep1
    m.call("evp_prep1", nx_block=NXB, ny_block=NYB, ilo=ilo, ihi=ihi, jlo=jlo, jhi=jhi, aice=S["aice"],
           vice=S["vice"], vsno=S["vsno"], tmask=tmask, strairxt=S["strairxT"], strairyt=S["strairyT"],
           strairx=O["strairx"], strairy=O["strairy"], tmass=O["tmass"], icetmask=icetmask)

    # to_ugrid works on (nx_block, ny_block, max_blocks) module arrays
    def as3(f):
        t = fe.FArr([(1, NXB), (1, NYB), (1, 1)])
        t.a[:, :, 0] = f.a
        return t

    ns.update(tarea=as3(F["tarea"]), uarea=as3(F["uarea"]))

    def to_grid(which, src):
        w2 = fe.FArr([(1, NXB), (1, NYB), (1, 1)])
        m.call(which, work1=as3(src), work2=w2)
        out = z()
        out.a[...] = w2.a[:, :, 0]
        return out

    O["umass"] = to_grid("to_ugrid", O["tmass"])                     # evp :218
    O["aiu"] = to_grid("to_ugrid", S["aice_init"])                   # :219
    O["strairx"] = to_grid("to_ugrid", O["strairx"])                 # :240 t2ugrid_vector (halo left out)
    O["strairy"] = to_grid("to_ugrid", O["strairy"])                 # :241
    n_list = NXB * NYB
    lists = {n: fe.FArr([(1, n_list)], dtype=int) for n in ("indxti", "indxtj", "indxui", "indxuj")}
    r = m.call("evp_prep2", nx_block=NXB, ny_block=NYB, ilo=ilo, ihi=ihi, jlo=jlo, jhi=jhi, icellt=0, icellu=0,
               aiu=O["aiu"], umass=O["umass"], umassdti=O["umassdti"], fcor=F["fcor"], umask=umask,
               uocn=S["uocn"], vocn=S["vocn"], strairx=O["strairx"], strairy=O["strairy"],
               ss_tltx=S["ss_tltx"], ss_tlty=S["ss_tlty"], icetmask=icetmask, iceumask=iceumask, fm=S["fm"],
               dt=DT, strtltx=O["strtltx"], strtlty=O["strtlty"], strocnx=O["strocnx"], strocny=O["strocny"],
               strintx=O["strintx"], strinty=O["strinty"], waterx=O["waterx"], watery=O["watery"],
               forcex=O["forcex"], forcey=O["forcey"], uvel_init=O["uvel_init"], vvel_init=O["vvel_init"],
               uvel=S["uvel"], vvel=S["vvel"], **lists, **ST)
    icellt, icellu = r["icellt"], r["icellu"]
    strtmp = fe.FArr([(1, NXB), (1, NYB), (1, 8)])
    for ksub in range(1, NDTE + 1):                                  # evp :336-410, halo left out
        m.call("stress", nx_block=NXB, ny_block=NYB, ksub=ksub, icellt=icellt, indxti=lists["indxti"],
               indxtj=lists["indxtj"], uvel=S["uvel"], vvel=S["vvel"], dxt=F["dxt"], dyt=F["dyt"],
               dxhy=F["dxhy"], dyhx=F["dyhx"], cxp=F["cxp"], cyp=F["cyp"], cxm=F["cxm"], cym=F["cym"],
               tarear=F["tarear"], tinyarea=F["tinyarea"], strength=S["strength"], shear=O["shear"],
               divu=O["divu"], prs_sig=O["prs_sig"], rdg_conv=O["rdg_conv"], rdg_shear=O["rdg_shear"],
               str=strtmp, **ST)
        m.call("stepu", nx_block=NXB, ny_block=NYB, icellu=icellu, cw=S["Cw"], indxui=lists["indxui"],
               indxuj=lists["indxuj"], aiu=O["aiu"], str=strtmp, uocn=S["uocn"], vocn=S["vocn"],
               waterx=O["waterx"], watery=O["watery"], forcex=O["forcex"], forcey=O["forcey"],
               umassdti=O["umassdti"], fm=S["fm"], uarear=F["uarear"], strocnx=O["strocnx"],
               strocny=O["strocny"], strintx=O["strintx"], strinty=O["strinty"], uvel_init=O["uvel_init"],
               vvel_init=O["vvel_init"], uvel=S["uvel"], vvel=S["vvel"])
    m.call("evp_finish", nx_block=NXB, ny_block=NYB, icellu=icellu, cw=S["Cw"], indxui=lists["indxui"],
           indxuj=lists["indxuj"], uvel=S["uvel"], vvel=S["vvel"], uocn=S["uocn"], vocn=S["vocn"], aiu=O["aiu"],
           fm=S["fm"], strintx=O["strintx"], strinty=O["strinty"], strairx=O["strairx"], strairy=O["strairy"],
           strocnx=O["strocnx"], strocny=O["strocny"], strocnxt=O["strocnxT"], strocnyt=O["strocnyT"])
    O["strocnxT"] = to_grid("to_tgrid", O["strocnxT"])               # :499-500 u2tgrid_vector (halo left out)
    O["strocnyT"] = to_grid("to_tgrid", O["strocnyT"])
    sig1, sig2 = z(), z()
    m.call("principal_stress", nx_block=NXB, ny_block=NYB, stressp_1=ST["stressp_1"], stressm_1=ST["stressm_1"],
           stress12_1=ST["stress12_1"], prs_sig=O["prs_sig"], sig1=sig1, sig2=sig2)
    res = {n: f.to_numpy() for n, f in O.items()}
    res.update({n: f.to_numpy() for n, f in ST.items()})
    res.update(uvel=S["uvel"].to_numpy(), vvel=S["vvel"].to_numpy(), fm=S["fm"].to_numpy(),
               iceumask=iceumask.to_numpy().astype(np.int32), icetmask=icetmask.to_numpy().astype(np.int32),
               sig1=sig1.to_numpy(), sig2=sig2.to_numpy(), icellt=np.int64(icellt), icellu=np.int64(icellu))
    for n in ("arlx1i", "brlx", "denom1", "revp", "ecci"):
        res["k_" + n] = np.float64(k[n])
    return res


def oracle_case(g, s):
    """The same block as a fixtures.Case for oracle.run_case(..., no_halo=True)."""
    from cice5_b200 import fixtures as fx
    dom = fx.Domain(NXB - 2, NYB - 2, ew_cyclic=True, ns_tripole=False, ns_open=True, grid_tripole=False,
                    name="ref_block")
    f = {k: np.ascontiguousarray(v) for k, v in g.items()}
    grid = fx.Grid(dom=dom, f=f)
    inp = {k: s[k].copy() for k in fx.IN_FIELDS}
    io = {k: s[k].copy() for k in fx.INOUT_FIELDS}
    io["iceumask"] = s["iceumask"].copy()
    return fx.Case(grid=grid, dt=DT, ndte=NDTE, inp=inp, io=io, out=fx.zeroed_outputs(dom))


COMPARE_IO = ("uvel", "vvel", "fm", "iceumask") + tuple(f"stress{k}_{c}" for k in ("p", "m", "12") for c in (1, 2, 3, 4))
COMPARE_OUT = ("strairx", "strairy", "strtltx", "strtlty", "strocnx", "strocny", "strintx", "strinty", "strocnxT",
               "strocnyT", "divu", "shear", "rdg_conv", "rdg_shear", "prs_sig", "uvel_init", "vvel_init", "tmass",
               "icetmask")


def run_reference_strength(m, s, cellmask, kstrength=1, krdg_partic=1, krdg_redist=1, mu_rdg=3.0, cf=17.0):
    """ice_strength (source/ice_mechred.F90:2111) from the reference text on the cells of `cellmask`
    (the icellt list of evp, source/ice_dyn_evp.F90:291-301), list order = the reference's (j outer)."""
    A = fe.FArr.from_numpy
    m.ns.update(ncat=NCAT, kstrength=kstrength, krdg_partic=krdg_partic, krdg_redist=krdg_redist, mu_rdg=mu_rdg,
                cf=cf)
    jj, ii = np.nonzero(cellmask)
    n_list = NXB * NYB
    indxi, indxj = fe.FArr([(1, n_list)], dtype=int), fe.FArr([(1, n_list)], dtype=int)
    indxi.a[:ii.size], indxj.a[:jj.size] = ii + 1, jj + 1

    def cat(a):                                   # (ncat, NYB, NXB) -> Fortran (nx_block, ny_block, ncat)
        t = fe.FArr([(1, NXB), (1, NYB), (1, NCAT)])
        t.a[...] = np.transpose(a, (2, 1, 0))
        return t

    strength = fe.FArr([(1, NXB), (1, NYB)])
    strength.fill(-7.0)                            # intent(out): the routine must overwrite everything
    m.call("ice_strength", nx_block=NXB, ny_block=NYB, ilo=2, ihi=NXB - 1, jlo=2, jhi=NYB - 1, icells=int(ii.size),
           indxi=indxi, indxj=indxj, aice=A(s["aice"]), vice=A(s["vice"]), aice0=A(s["aice0"]),
           aicen=cat(s["aicen"]), vicen=cat(s["vicen"]), strength=strength)
    return strength.to_numpy()


# ---------------------------------------------------------------------------------------------
# variants: CPP macros of the reference build, module variables, and the oracle configuration that
# must reproduce each bit for bit
# ---------------------------------------------------------------------------------------------
_ANGLE = 25.0 * np.pi / 180.0
VARIANTS = {
    # name: (CPP macros, module overrides, revised_evp, oracle cfg keywords)
    "default": ((), {}, False, {}),
    "coupled": (("coupled",), {}, False, {"tilt_mode": 1}),
    "auscom_slope": (("AusCOM", "coupled"), {"cosw": float(np.cos(_ANGLE)), "sinw": float(np.sin(_ANGLE)),
                                             "use_ocnslope": True, "dragio": 0.00536},
                     False, {"tilt_mode": 1, "cosw": float(np.cos(_ANGLE)), "sinw": float(np.sin(_ANGLE))}),
    "auscom_noslope": (("AusCOM", "coupled"), {"cosw": float(np.cos(_ANGLE)), "sinw": float(np.sin(_ANGLE)),
                                               "use_ocnslope": False, "dragio": 0.00536},
                       False, {"tilt_mode": 0, "cosw": float(np.cos(_ANGLE)), "sinw": float(np.sin(_ANGLE))}),
    "revised": ((), {}, True, {"revised_evp": True}),
}
STRENGTH_VARIANTS = ((1, 1, 1), (1, 0, 0), (1, 0, 1), (1, 1, 0), (0, 1, 1))   # kstrength, krdg_partic, krdg_redist
SEEDS = (1, 2, 3)
IN_GRID = ("dxt", "dyt", "dxhy", "dyhx", "cxp", "cyp", "cxm", "cym", "tarea", "uarea", "tarear", "uarear",
           "tinyarea", "fcor", "tmask", "umask")
OUT_ALL = COMPARE_IO + COMPARE_OUT + ("sig1", "sig2")


def main(out_path):
    pack = {}
    for vname, (macros, overrides, revised, _okw) in VARIANTS.items():
        m = load_modules(macros, overrides)
        for seed in SEEDS:
            g, s = random_block(seed)
            res = run_reference_sequence(m, g, s, revised_evp=revised)
            key = f"{vname}__{seed}__"
            if vname == "default":                 # inputs are per seed, shared by the variants
                for n in IN_GRID:
                    pack[f"in__{seed}__{n}"] = g[n]
                for n, v in s.items():
                    pack[f"in__{seed}__{n}"] = v
                mask = (res["icetmask"] == 1).astype(np.int32)
                for ks, kp, kr in STRENGTH_VARIANTS:
                    pack[f"strength__{seed}__{ks}{kp}{kr}"] = run_reference_strength(m, s, mask, ks, kp, kr)
            for n in OUT_ALL:
                pack[key + n] = res[n]
            pack[key + "lists"] = np.array([res["icellt"], res["icellu"]], np.int64)
            pack[key + "consts"] = np.array([res["k_" + n] for n in ("arlx1i", "brlx", "denom1", "revp", "ecci")])
    pack["meta"] = np.array([NXB, NYB, NDTE, NCAT], np.int64)
    pack["dt"] = np.float64(DT)
    np.savez_compressed(out_path, **pack)
    print(f"{out_path}: {len(pack)} arrays, {os.path.getsize(out_path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_block_vectors.npz"))
