#!/usr/bin/env python
"""Golden vectors of the EAP dynamics from the REFERENCE'S OWN TEXT (run in the build container only).

source/ice_dyn_eap.F90 executed mechanically by tests/golden/fortran_exec.py:

  unit level   w1, w2, s11kr ... s22ks at random arguments; init_eap's table loop (:567-618) on a subset of table
               entries -- the one edit of the text is the three loop headers `do ia=1,na_yield` etc., which get the
               subset's bounds; update_stress_rdg and calc_ffrac on random arguments
  block level  the driver's call order (eap, :150-482) on one small block without halo updates: evp_prep1, to_ugrid,
               evp_prep2, the structure-tensor reset where icetmask == 0 (:274-289, restated here: it is inline driver
               text), then ndte x (stress_eap, stepu, stepa every 10th subcycle), evp_finish, to_tgrid

Intrinsics.  init_eap's table code runs with the C library's sin / cos / tan / exp (python's math module), as does
oracle/eap_oracle.inc.  update_stress_rdg / calc_ffrac / stepa / stress_eap run with atan2, sin, cos bound to
cice5_b200/csrc/evp_trig.h (through the oracle library's evp_oracle_sin / cos / atan2): the one definition the device
and the oracle share -- a Fortran processor is free in the last bit of these intrinsics, and the table look-up
(`int((x-piq-pi)*invdx)`) turns a last-bit difference into another table entry.

    python tests/golden/make_eap_vectors.py        # rewrites tests/golden/ref_eap_vectors.npz
"""
from __future__ import annotations

import math
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import fortran_exec as fe  # noqa: E402
import make_reference_vectors as mrv  # noqa: E402
from oracle import oracle as orc  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(HERE, "ref_eap_vectors.npz")
NXB, NYB = mrv.NXB, mrv.NYB
NDTE, DT = 12, 3600.0          # stepa runs on subcycles 1 and 11
TABLE_SUBSET = {"ia": (1, 10, 20, 21), "ix": (1, 7, 23, 41), "iy": (1, 12, 30, 41)}
EAP_FIELDS = orc.EAP_CORNER + orc.EAP_AVG + orc.EAP_OUT


def eap_lines(defined=()):
    return fe.preprocess(open(os.path.join(REF, "source/ice_dyn_eap.F90")).read(), defined)


def load_eap(shared_trig: bool):
    """The EVP module of make_reference_vectors plus the EAP routines; intrinsics as the docstring says."""
    m = mrv.load_modules()
    lines = eap_lines()
    fe.parse_parameters(lines, m.ns)                       # nx_yield, ny_yield, na_yield
    if shared_trig:
        m.ns.update(sin=orc.trig("sin"), cos=orc.trig("cos"), atan2=orc.trig("atan2"))
    else:
        m.ns.update(sin=math.sin, cos=math.cos, atan2=math.atan2)
    m.ns.update(tan=math.tan)
    m.load(lines, ["w1", "w2", "s11kr", "s12kr", "s22kr", "s11ks", "s12ks", "s22ks"])
    m.load(lines, ["calc_ffrac", "update_stress_rdg", "stepa", "stress_eap"])
    return m, lines


def table_farr(t):
    """(na, ny, nx) C array -> Fortran (nx_yield, ny_yield, na_yield)."""
    f = fe.FArr([(1, 41), (1, 41), (1, 21)])
    f.a[...] = np.transpose(t, (2, 1, 0))
    return f


def unit_vectors(pack):
    r = np.random.default_rng(11)
    m, lines = load_eap(shared_trig=False)
    # w1, w2, the six table kernels
    a = r.uniform(0.5, 1.0, 64)
    pack["unit__w_a"] = a
    pack["unit__w1"] = np.array([m.call("w1", a=float(v)) for v in a])
    pack["unit__w2"] = np.array([m.call("w2", a=float(v)) for v in a])
    xyz = np.stack([r.uniform(math.pi / 4, 9 * math.pi / 4, 200), r.uniform(0.0, math.pi, 200),
                    r.uniform(-math.pi / 2, math.pi / 2, 200)], axis=1)
    pack["unit__kernel_xyz"] = xyz
    phi = math.pi / 12.0
    for k, name in enumerate(("s11kr", "s12kr", "s22kr", "s11ks", "s12ks", "s22ks")):
        pack[f"unit__kernel_{k}"] = np.array([m.call(name, x=float(x), y=float(y), z=float(z), phi=phi) for x, y, z in xyz])
    # init_eap's table loop on a subset: the reference's statements, only the loop headers narrowed
    sub = eap_lines()
    for var, hi in (("ia", "na_yield"), ("ix", "nx_yield"), ("iy", "ny_yield")):
        n = sum(1 for l in sub if l.replace(" ", "") == f"do{var}=1,{hi}")
        assert n == 1, (var, n)
        sub = [f"do {var} = {var}_lo, {var}_hi" if l.replace(" ", "") == f"do{var}=1,{hi}" else l for l in sub]
    tabs = {n: fe.FArr([(1, 41), (1, 41), (1, 21)]) for n in orc.EAP_TABLES}
    z3 = lambda: fe.FArr([(1, NXB), (1, NYB), (1, 1)])
    m.ns.update(tabs)
    m.ns.update({n: z3() for n in ("e11", "e12", "e22", "s11", "s12", "s22", "yieldstress11", "yieldstress12",
                                   "yieldstress22", "a11_1", "a11_2", "a11_3", "a11_4", "a12_1", "a12_2", "a12_3",
                                   "a12_4")})
    m.ns.update(nblocks=1, nx_block=NXB, ny_block=NYB, init_evp=lambda *a, **k: None)
    m.load(sub, ["init_eap"])
    entries = []
    for ia in TABLE_SUBSET["ia"]:
        for ix in TABLE_SUBSET["ix"]:
            for iy in TABLE_SUBSET["iy"]:
                m.ns.update(ia_lo=ia, ia_hi=ia, ix_lo=ix, ix_hi=ix, iy_lo=iy, iy_hi=iy)
                m.call("init_eap", dt=DT)
                entries.append((ia, ix, iy))
    pack["unit__table_entries"] = np.array(entries, np.int64)
    for n in orc.EAP_TABLES:
        pack[f"unit__table_{n}"] = np.array([tabs[n](ix, iy, ia) for ia, ix, iy in entries])
    assert float(np.asarray(m.ns["a11_1"].a).min()) == 0.5           # init_eap's state reset ran too

    # update_stress_rdg, calc_ffrac with the shared intrinsics and the oracle's (C-generated, pinned above) tables
    m, _ = load_eap(shared_trig=True)
    T = orc.eap_tables()
    m.ns.update({n: table_farr(T[n]) for n in orc.EAP_TABLES})
    n = 400
    arg = {"divu": r.normal(0, 3e3, n), "tension": r.normal(0, 3e3, n), "shear": r.normal(0, 3e3, n),
           "a11": r.uniform(0.02, 0.98, n), "strength": r.uniform(1e3, 5e4, n)}
    # a valid structure tensor: eigenvalues in [0, 1], i.e. a12^2 <= a11 a22 (else ka leaves the table, :1601)
    arg["a12"] = r.uniform(-0.98, 0.98, n) * np.sqrt(arg["a11"] * (1.0 - arg["a11"]))
    arg["divu"][:8] = 0.0; arg["tension"][:8] = 0.0; arg["shear"][:8] = 0.0        # x = 0 branch
    arg["a11"][8:16] = 0.5; arg["a12"][8:16] = 0.0                                # atan2(0, 0)
    res = np.zeros((n, 2, 5))
    for i in range(n):
        for last in (0, 1):
            o = m.call("update_stress_rdg", ksub=12 if last else 3, ndte=12, stressp=0.0, stressm=0.0, stress12=0.0,
                       alphar=-7.0, alphas=-7.0, **{k: float(v[i]) for k, v in arg.items()})
            res[i, last] = [o["stressp"], o["stressm"], o["stress12"], o["alphar"], o["alphas"]]
    for k, v in arg.items():
        pack[f"unit__rdg_{k}"] = v
    pack["unit__rdg_out"] = res
    fa = {"stressp": r.normal(0, 2e3, n), "stressm": r.normal(0, 2e3, n), "stress12": r.normal(0, 2e3, n),
          "a1x": r.uniform(-0.5, 1.0, n)}
    fa["stressp"][:40] = -np.abs(fa["stressp"][:40]) * 3                            # compressive states
    fa["stressp"][40:60] = np.abs(fa["stressp"][40:60]) * 3                         # divergent states
    out = np.zeros((n, 2))
    for i in range(n):
        for b in (1, 2):
            o = m.call("calc_ffrac", blockno=b, mresult=0.0, **{k: float(v[i]) for k, v in fa.items()})
            out[i, b - 1] = o["mresult"]
    for k, v in fa.items():
        pack[f"unit__ffrac_{k}"] = v
    pack["unit__ffrac_out"] = out


def eap_block_state(seed):
    r = np.random.default_rng(1000 + seed)
    sh = (NYB, NXB)
    e = {}
    for c in (1, 2, 3, 4):
        e[f"a11_{c}"] = r.uniform(0.3, 0.9, sh)
        e[f"a12_{c}"] = r.uniform(-0.9, 0.9, sh) * np.sqrt(e[f"a11_{c}"] * (1.0 - e[f"a11_{c}"]))
    e["a11"], e["a12"] = r.uniform(0.3, 0.9, sh), r.uniform(-0.3, 0.3, sh)
    for n in orc.EAP_OUT:
        e[n] = r.uniform(-1.0, 1.0, sh)              # the driver zeroes them (:166-182)
    return e


def run_reference_sequence_eap(m, g, s, e):
    """eap's call order (source/ice_dyn_eap.F90:150-482) on one block, halos left out."""
    A = fe.FArr.from_numpy
    ns = m.ns
    ilo, ihi, jlo, jhi = 2, NXB - 1, 2, NYB - 1
    blk = types.SimpleNamespace(ilo=ilo, ihi=ihi, jlo=jlo, jhi=jhi)
    F = {k: A(v) for k, v in g.items() if k not in ("tmask", "umask")}
    tmask, umask = A(g["tmask"] > 0), A(g["umask"] > 0)
    ns.update(ndte=NDTE, revised_evp=False, my_task=0, master_task=0, distrb_info=None,
              dxt=F["dxt"], dyt=F["dyt"], tmask=tmask, nblocks=1, blocks_ice=lambda i: i, get_block=lambda *_: blk,
              global_minval=lambda arr, _d, msk: float(arr.a[1:-1, 1:-1][msk.a[1:-1, 1:-1]].min()))
    m.call("set_evp_parameters", dt=DT)
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
    E = {n: A(e[n]) for n in EAP_FIELDS}
    for n in orc.EAP_OUT:                                            # :166-182
        E[n].fill(0.0)
    icetmask = zi()
    m.call("evp_prep1", nx_block=NXB, ny_block=NYB, ilo=ilo, ihi=ihi, jlo=jlo, jhi=jhi, aice=S["aice"],
           vice=S["vice"], vsno=S["vsno"], tmask=tmask, strairxt=S["strairxT"], strairyt=S["strairyT"],
           strairx=O["strairx"], strairy=O["strairy"], tmass=O["tmass"], icetmask=icetmask)

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

    O["umass"] = to_grid("to_ugrid", O["tmass"])
    O["aiu"] = to_grid("to_ugrid", S["aice_init"])
    O["strairx"] = to_grid("to_ugrid", O["strairx"])
    O["strairy"] = to_grid("to_ugrid", O["strairy"])
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
    # :274-289, inline driver text: a11_k = p5, a12_k = c0 where icetmask == 0, every cell of the block
    off = icetmask.a == 0
    for c in (1, 2, 3, 4):
        E[f"a11_{c}"].a[off] = 0.5
        E[f"a12_{c}"].a[off] = 0.0
    strtmp = fe.FArr([(1, NXB), (1, NYB), (1, 8)])
    acorn = {n: E[n] for n in orc.EAP_CORNER}
    for ksub in range(1, NDTE + 1):
        m.call("stress_eap", nx_block=NXB, ny_block=NYB, ksub=ksub, ndte=NDTE, icellt=icellt,
               indxti=lists["indxti"], indxtj=lists["indxtj"], arlx1i=ns["arlx1i"], denom1=ns["denom1"],
               uvel=S["uvel"], vvel=S["vvel"], dxt=F["dxt"], dyt=F["dyt"], dxhy=F["dxhy"], dyhx=F["dyhx"],
               cxp=F["cxp"], cyp=F["cyp"], cxm=F["cxm"], cym=F["cym"], tarear=F["tarear"], strength=S["strength"],
               a11=E["a11"], a12=E["a12"], shear=O["shear"], divu=O["divu"], e11=E["e11"], e12=E["e12"],
               e22=E["e22"], s11=E["s11"], s12=E["s12"], s22=E["s22"], yieldstress11=E["yieldstress11"],
               yieldstress12=E["yieldstress12"], yieldstress22=E["yieldstress22"], prs_sig=O["prs_sig"],
               rdg_conv=O["rdg_conv"], rdg_shear=O["rdg_shear"], strtmp=strtmp, **acorn, **ST)
        m.call("stepu", nx_block=NXB, ny_block=NYB, icellu=icellu, cw=S["Cw"], indxui=lists["indxui"],
               indxuj=lists["indxuj"], aiu=O["aiu"], str=strtmp, uocn=S["uocn"], vocn=S["vocn"],
               waterx=O["waterx"], watery=O["watery"], forcex=O["forcex"], forcey=O["forcey"],
               umassdti=O["umassdti"], fm=S["fm"], uarear=F["uarear"], strocnx=O["strocnx"],
               strocny=O["strocny"], strintx=O["strintx"], strinty=O["strinty"], uvel_init=O["uvel_init"],
               vvel_init=O["vvel_init"], uvel=S["uvel"], vvel=S["vvel"])
        if ksub % 10 == 1:                                           # :408
            m.call("stepa", nx_block=NXB, ny_block=NYB, dtei=ns["dtei"], icellt=icellt, indxti=lists["indxti"],
                   indxtj=lists["indxtj"], a11=E["a11"], a12=E["a12"], **acorn, **ST)
    m.call("evp_finish", nx_block=NXB, ny_block=NYB, icellu=icellu, cw=S["Cw"], indxui=lists["indxui"],
           indxuj=lists["indxuj"], uvel=S["uvel"], vvel=S["vvel"], uocn=S["uocn"], vocn=S["vocn"], aiu=O["aiu"],
           fm=S["fm"], strintx=O["strintx"], strinty=O["strinty"], strairx=O["strairx"], strairy=O["strairy"],
           strocnx=O["strocnx"], strocny=O["strocny"], strocnxt=O["strocnxT"], strocnyt=O["strocnyT"])
    O["strocnxT"] = to_grid("to_tgrid", O["strocnxT"])
    O["strocnyT"] = to_grid("to_tgrid", O["strocnyT"])
    res = {n: f.to_numpy() for n, f in O.items()}
    res.update({n: f.to_numpy() for n, f in ST.items()})
    res.update({n: f.to_numpy() for n, f in E.items()})
    res.update(uvel=S["uvel"].to_numpy(), vvel=S["vvel"].to_numpy(), fm=S["fm"].to_numpy(),
               iceumask=iceumask.to_numpy().astype(np.int32), icetmask=icetmask.to_numpy().astype(np.int32),
               icellt=np.int64(icellt), icellu=np.int64(icellu))
    return res


SEEDS = (1, 2)
OUT_ALL = mrv.COMPARE_IO + mrv.COMPARE_OUT + EAP_FIELDS


def main(out_path):
    pack = {}
    unit_vectors(pack)
    m, _ = load_eap(shared_trig=True)
    T = orc.eap_tables()
    m.ns.update({n: table_farr(T[n]) for n in orc.EAP_TABLES})
    for seed in SEEDS:
        g, s = mrv.random_block(seed)
        e = eap_block_state(seed)
        for n, v in e.items():
            pack[f"in__{seed}__{n}"] = v
        res = run_reference_sequence_eap(m, g, s, e)
        for n in OUT_ALL:
            pack[f"block__{seed}__{n}"] = res[n]
        pack[f"block__{seed}__lists"] = np.array([res["icellt"], res["icellu"]], np.int64)
    pack["meta"] = np.array([NXB, NYB, NDTE], np.int64)
    pack["dt"] = np.float64(DT)
    np.savez_compressed(out_path, **pack)
    print(f"{out_path}: {len(pack)} arrays, {os.path.getsize(out_path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main(OUT)
