#!/usr/bin/env python
"""Golden vectors for the first stage of horizontal transport from the REFERENCE'S OWN TEXT
(run in the build container only): `departure_points` of source/ice_transport_remap.F90:1493-1670, executed
mechanically by tests/golden/fortran_exec.py on seeded blocks -- first-order and midpoint-corrected departure points,
cells at rest, and a block whose velocities send a departure point more than one cell away (l_stop, istop, jstop).

    python tests/golden/make_remap_vectors.py        # rewrites tests/golden/ref_remap_vectors.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import fortran_exec as fe  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(HERE, "ref_remap_vectors.npz")
NXB, NYB = 14, 11            # block incl. the ghost ring


def load():
    def lines(rel):
        return fe.preprocess(open(os.path.join(REF, rel)).read(), ())
    ns = {}
    fe.parse_parameters(lines("drivers/cice/ice_constants.F90"), ns)
    ns.update(nu_diag=6, my_task=0)
    m = fe.Module(ns)
    m.load(lines("source/ice_transport_remap.F90"), ["departure_points"])
    return m


# ---- second stage: make_masks + construct_fields (+ limited_gradient), :867-1484 ------------------------------------
NCAT = 2
TRACER_TYPE = [1, 1, 1, 2, 2, 3]          # hice, hsno, Tsfc-like, qice-like (on hice), qsno-like (on hsno), a type 3
DEPEND = [0, 0, 0, 1, 2, 4]


def load_fields():
    def lines(rel):
        return fe.preprocess(open(os.path.join(REF, rel)).read(), ())
    ns = {}
    fe.parse_parameters(lines("drivers/cice/ice_constants.F90"), ns)
    ns.update(nu_diag=6, my_task=0, ncat=NCAT)
    m = fe.Module(ns)
    m.load(lines("source/ice_transport_remap.F90"), ["make_masks", "construct_fields", "limited_gradient"])
    return m


def has_dependents(depend):
    """init_transport, source/ice_transport_driver.F90:112-118"""
    h = [False] * len(depend)
    for d in depend:
        if d > 0:
            h[d - 1] = True
    return h


def fields_block(seed, open_frac):
    r = np.random.default_rng(seed)
    sh = (NYB, NXB)
    nt = len(TRACER_TYPE)
    hm = (r.uniform(0, 1, sh) < 0.85).astype(float)
    mm = r.uniform(0.05, 1.0, (NCAT + 1,) + sh)
    mm[r.uniform(0, 1, mm.shape) < open_frac] = 0.0            # no ice of this category
    mm[r.uniform(0, 1, mm.shape) < 0.05] = 5.0e-12             # below puny
    mm *= hm                                                   # no ice on land
    tm = r.uniform(0.1, 2.0, (NCAT, nt) + sh)
    tm[:, 2] = -r.uniform(0.5, 20.0, (NCAT,) + sh)             # a temperature: negative values
    tm[r.uniform(0, 1, tm.shape) < 0.12] = 0.0                 # tracer absent (snow-free ice ...)
    tm[r.uniform(0, 1, tm.shape) < 0.03] = 3.0e-12
    flat = r.uniform(0, 1, sh) < 0.15                          # locally uniform fields: zero gradients, 0/0 guards
    tm[:, :, flat] = 0.7
    return {"hm": hm, "xav": r.uniform(-0.04, 0.04, sh), "yav": r.uniform(-0.04, 0.04, sh),
            "xxav": r.uniform(0.07, 0.10, sh), "yyav": r.uniform(0.07, 0.10, sh), "mm": mm, "tm": tm}


def run_fields(m, d):
    """make_masks, then construct_fields for the open-water category and each ice category, as horizontal_remap calls
    them (:462-511).  Arrays: mm (ncat+1, NYB, NXB) = Fortran (nx_block, ny_block, 0:ncat); tm (ncat, ntrace, NYB, NXB) =
    Fortran (nx_block, ny_block, ntrace, ncat).  Outputs start with garbage."""
    nt = len(TRACER_TYPE)
    hd = has_dependents(DEPEND)
    geo = dict(nx_block=NXB, ny_block=NYB, ilo=2, ihi=NXB - 1, jlo=2, jhi=NYB - 1, nghost=1, ntrace=nt)
    icells = fe.FArr.from_numpy(np.zeros(NCAT + 1, dtype=np.int64), los=[0])
    indxi = fe.FArr.from_numpy(np.zeros((NCAT + 1, NXB * NYB), dtype=np.int64), los=[1, 0])
    indxj = fe.FArr.from_numpy(np.zeros((NCAT + 1, NXB * NYB), dtype=np.int64), los=[1, 0])
    mm = fe.FArr.from_numpy(d["mm"], los=[1, 1, 0])
    tm = fe.FArr.from_numpy(d["tm"])
    mmask = fe.FArr.from_numpy(np.full((NCAT + 1, NYB, NXB), -7.0), los=[1, 1, 0])
    tmask = fe.FArr.from_numpy(np.full((NCAT, nt, NYB, NXB), -7.0))
    hdf = fe.FArr.from_numpy(np.array(hd))
    m.call("make_masks", has_dependents=hdf, icells=icells, indxi=indxi, indxj=indxj, mm=mm, mmask=mmask, tm=tm,
           tmask=tmask, **geo)
    g = {k: fe.FArr.from_numpy(d[k]) for k in ("hm", "xav", "yav", "xxav", "yyav")}
    mc, mx, my = (fe.FArr.from_numpy(np.full((NCAT + 1, NYB, NXB), -7.0), los=[1, 1, 0]) for _ in range(3))
    tc, tx, ty = (fe.FArr.from_numpy(np.full((NCAT, nt, NYB, NXB), -7.0)) for _ in range(3))
    tt, dep = fe.FArr.from_numpy(np.array(TRACER_TYPE)), fe.FArr.from_numpy(np.array(DEPEND))
    S = slice(None)
    m.call("construct_fields", tracer_type=tt, depend=dep, has_dependents=hdf, icells=icells(0), indxi=indxi(S, 0),
           indxj=indxj(S, 0), mm=mm(S, S, 0), mc=mc(S, S, 0), mx=mx(S, S, 0), my=my(S, S, 0), mmask=mmask(S, S, 0),
           **g, **geo)
    for n in range(1, NCAT + 1):
        m.call("construct_fields", tracer_type=tt, depend=dep, has_dependents=hdf, icells=icells(n),
               indxi=indxi(S, n), indxj=indxj(S, n), mm=mm(S, S, n), mc=mc(S, S, n), mx=mx(S, S, n), my=my(S, S, n),
               mmask=mmask(S, S, n), tm=tm(S, S, S, n), tc=tc(S, S, S, n), tx=tx(S, S, S, n), ty=ty(S, S, S, n),
               tmask=tmask(S, S, S, n), **g, **geo)
    out = {k: v.to_numpy() for k, v in (("mc", mc), ("mx", mx), ("my", my), ("tc", tc), ("tx", tx), ("ty", ty),
                                         ("mmask", mmask), ("tmask", tmask))}
    out["icells"] = icells.to_numpy()
    return out


def block(seed, speed):
    r = np.random.default_rng(seed)
    sh = (NYB, NXB)
    d = {"uvel": r.uniform(-speed, speed, sh), "vvel": r.uniform(-speed, speed, sh),
         "dxu": r.uniform(2.0e4, 3.0e4, sh), "dyu": r.uniform(2.0e4, 3.0e4, sh),
         "HTN": r.uniform(2.0e4, 3.0e4, sh), "HTE": r.uniform(2.0e4, 3.0e4, sh)}
    rest = r.uniform(0, 1, sh) < 0.2                     # cells at rest: the midpoint branch is skipped
    d["uvel"][rest] = 0.0
    d["vvel"][rest] = 0.0
    half = r.uniform(0, 1, sh) < 0.1                     # one component zero
    d["vvel"][half] = 0.0
    return d


def run(m, d, dt, midpt):
    """One call; arrays are (NYB, NXB) C order == Fortran (nx_block, ny_block).  The out arrays start with garbage:
    the routine zeroes them itself (:1557-1558)."""
    a = {k.lower(): fe.FArr.from_numpy(v) for k, v in d.items()}
    dpx, dpy = fe.FArr.from_numpy(np.full((NYB, NXB), -7.0)), fe.FArr.from_numpy(np.full((NYB, NXB), -7.0))
    r = m.call("departure_points", nx_block=NXB, ny_block=NYB, ilo=2, ihi=NXB - 1, jlo=2, jhi=NYB - 1, nghost=1, dt=dt,
               dpx=dpx, dpy=dpy, l_dp_midpt=midpt, l_stop=False, istop=0, jstop=0, **a)
    return {"dpx": dpx.to_numpy(), "dpy": dpy.to_numpy(), "l_stop": int(bool(r["l_stop"])), "istop": int(r["istop"]),
            "jstop": int(r["jstop"])}


# ---- third stage: locate_triangles, triangle_coordinates, transport_integrals, update_fields (:1680-3729) -----------
NG, NV = 6, 3


def load_fluxes():
    def lines(rel):
        return fe.preprocess(open(os.path.join(REF, rel)).read(), ())
    ns = {}
    fe.parse_parameters(lines("drivers/cice/ice_constants.F90"), ns)
    ns.update(nu_diag=6, my_task=0, ncat=NCAT, ngroups=NG, nvert=NV, char_len=80, bugcheck=False,
              p5625m=-9.0 / 16.0, p52083=25.0 / 48.0)
    m = fe.Module(ns)
    m.load(lines("source/ice_transport_remap.F90"),
           ["make_masks", "construct_fields", "limited_gradient", "departure_points", "locate_triangles",
            "triangle_coordinates", "transport_integrals", "update_fields"])
    return m


def run_triangles(m, edge, dpx, dpy, dxu, dyu):
    """locate_triangles + triangle_coordinates (integral_order = 3) for one edge, as horizontal_remap calls them
    (:668-690, :719-741).  Returns the arrays in this repository's order: xp, yp (6, 4, NYB, NXB), ..."""
    xp, yp = (fe.FArr.from_numpy(np.full((NG, NV + 1, NYB, NXB), -7.0), los=[1, 1, 0, 1]) for _ in range(2))
    iflux, jflux = (fe.FArr.from_numpy(np.zeros((NG, NYB, NXB), dtype=np.int64)) for _ in range(2))
    triarea = fe.FArr.from_numpy(np.full((NG, NYB, NXB), -7.0))
    icells = fe.FArr.from_numpy(np.zeros(NG, dtype=np.int64))
    indxi, indxj = (fe.FArr.from_numpy(np.zeros((NG, NXB * NYB), dtype=np.int64)) for _ in range(2))
    edgearea = fe.FArr.from_numpy(np.zeros((NYB, NXB)))
    geo = dict(nx_block=NXB, ny_block=NYB)
    m.call("locate_triangles", ilo=2, ihi=NXB - 1, jlo=2, jhi=NYB - 1, nghost=1, edge=edge, icells=icells, indxi=indxi,
           indxj=indxj, dpx=fe.FArr.from_numpy(dpx), dpy=fe.FArr.from_numpy(dpy), dxu=fe.FArr.from_numpy(dxu),
           dyu=fe.FArr.from_numpy(dyu), xp=xp, yp=yp, iflux=iflux, jflux=jflux, triarea=triarea, l_fixed_area=False,
           edgearea=edgearea, **geo)
    m.call("triangle_coordinates", integral_order=3, icells=icells, indxi=indxi, indxj=indxj, xp=xp, yp=yp, **geo)
    handles = dict(xp=xp, yp=yp, iflux=iflux, jflux=jflux, triarea=triarea, icells=icells, indxi=indxi, indxj=indxj)
    out = {"xp": xp.to_numpy(), "yp": yp.to_numpy(), "iflux": iflux.to_numpy().astype(np.int32),
           "jflux": jflux.to_numpy().astype(np.int32), "triarea": triarea.to_numpy(),
           "icells": icells.to_numpy().astype(np.int32), "edgearea": edgearea.to_numpy()}
    return handles, out


def triangle_block(seed, amp, shear):
    """Corner displacements up to `amp` cells in every direction; `shear` adds a gradient along the edges so that the
    two ends of an edge leave on opposite sides (the xic branches); a fifth of the corners at rest."""
    r = np.random.default_rng(seed)
    sh = (NYB, NXB)
    dxu, dyu = r.uniform(2.0e4, 3.0e4, sh), r.uniform(2.0e4, 3.0e4, sh)
    ii = np.arange(NXB)[None, :] + np.arange(NYB)[:, None]
    sx = r.uniform(-amp, amp, sh) + shear * np.where(ii % 2 == 0, 1.0, -1.0) * r.uniform(0.0, amp, sh)
    sy = r.uniform(-amp, amp, sh) + shear * np.where(ii % 2 == 0, -1.0, 1.0) * r.uniform(0.0, amp, sh)
    rest = r.uniform(0, 1, sh) < 0.2
    sx[rest] = 0.0
    sy[rest] = 0.0
    sy[r.uniform(0, 1, sh) < 0.05] = 0.0                    # motion along the edge only
    return {"dpx": sx * dxu, "dpy": sy * dyu, "dxu": dxu, "dyu": dyu}


def run_pipeline(m, d, vel, dt, tarear):
    """horizontal_remap (:452-850) for one block without its ghost-cell updates: the reference's routines in the
    reference's order, on the inputs of a construct_fields case + velocities."""
    nt = len(TRACER_TYPE)
    hd = has_dependents(DEPEND)
    geo = dict(nx_block=NXB, ny_block=NYB)
    blk = dict(ilo=2, ihi=NXB - 1, jlo=2, jhi=NYB - 1, **geo)
    rec = run_fields(m, d)
    F = fe.FArr.from_numpy
    dpx, dpy = F(np.zeros((NYB, NXB))), F(np.zeros((NYB, NXB)))
    r = m.call("departure_points", nghost=1, dt=dt, dpx=dpx, dpy=dpy, l_dp_midpt=True, l_stop=False, istop=0, jstop=0,
               uvel=F(vel["uvel"]), vvel=F(vel["vvel"]), dxu=F(vel["dxu"]), dyu=F(vel["dyu"]), htn=F(vel["HTN"]),
               hte=F(vel["HTE"]), **blk)
    assert not r["l_stop"]
    tt, dep = F(np.array(TRACER_TYPE)), F(np.array(DEPEND))
    mc, mx, my = (F(rec[k], los=[1, 1, 0]) for k in ("mc", "mx", "my"))
    tc, tx, ty = (F(rec[k]) for k in ("tc", "tx", "ty"))
    S = slice(None)
    out = {"dpx": dpx.to_numpy(), "dpy": dpy.to_numpy()}
    flux = {}
    for edge in ("east", "north"):
        h, tri = run_triangles(m, edge, dpx.to_numpy(), dpy.to_numpy(), vel["dxu"], vel["dyu"])
        for k, v in tri.items():
            out[f"{edge}/{k}"] = v
        mflx = F(np.full((NCAT + 1, NYB, NXB), -7.0), los=[1, 1, 0])
        mtflx = F(np.full((NCAT, nt, NYB, NXB), -7.0))
        common = dict(ntrace=nt, icells=h["icells"], indxi=h["indxi"], indxj=h["indxj"], tracer_type=tt, depend=dep,
                      integral_order=3, triarea=h["triarea"], iflux=h["iflux"], jflux=h["jflux"], xp=h["xp"], yp=h["yp"], **geo)
        m.call("transport_integrals", mc=mc(S, S, 0), mx=mx(S, S, 0), my=my(S, S, 0), mflx=mflx(S, S, 0), **common)
        for n in range(1, NCAT + 1):
            m.call("transport_integrals", mc=mc(S, S, n), mx=mx(S, S, n), my=my(S, S, n), mflx=mflx(S, S, n),
                   tc=tc(S, S, S, n), tx=tx(S, S, S, n), ty=ty(S, S, S, n), mtflx=mtflx(S, S, S, n), **common)
        flux[edge] = (mflx, mtflx)
        out[f"{edge}/mflx"], out[f"{edge}/mtflx"] = mflx.to_numpy(), mtflx.to_numpy()
    mm, tm = F(d["mm"], los=[1, 1, 0]), F(d["tm"])
    stops = []
    common = dict(ntrace=nt, tracer_type=tt, depend=dep, tarear=F(tarear), **blk)
    r = m.call("update_fields", l_stop=False, istop=0, jstop=0, mflxe=flux["east"][0](S, S, 0), mflxn=flux["north"][0](S, S, 0),
               mm=mm(S, S, 0), **common)
    stops.append((bool(r["l_stop"]), int(r["istop"]), int(r["jstop"])))
    for n in range(1, NCAT + 1):
        r = m.call("update_fields", l_stop=False, istop=0, jstop=0, mflxe=flux["east"][0](S, S, n),
                   mflxn=flux["north"][0](S, S, n), mm=mm(S, S, n), mtflxe=flux["east"][1](S, S, S, n),
                   mtflxn=flux["north"][1](S, S, S, n), tm=tm(S, S, S, n), **common)
        stops.append((bool(r["l_stop"]), int(r["istop"]), int(r["jstop"])))
    out["mm_new"], out["tm_new"] = mm.to_numpy(), tm.to_numpy()
    out["stops"] = np.array([[int(a), b, c] for a, b, c in stops], dtype=np.int32)
    return out


# ---- the caller's conversions: state_to_tracers / tracers_to_state (source/ice_transport_driver.F90:789-1002) ---------
NTRCR, NT_QSNO, NSLYR = 7, 6, 1          # Tsfc, 4 x qice, qsno, one more; nt_qsno 1-based


def load_driver():
    def lines(rel):
        return fe.preprocess(open(os.path.join(REF, rel)).read(), ())
    ns = {}
    fe.parse_parameters(lines("drivers/cice/ice_constants.F90"), ns)
    ns.update(nu_diag=6, my_task=0, ncat=NCAT, nslyr=NSLYR, nt_qsno=NT_QSNO)
    m = fe.Module(ns)
    m.load(lines("source/ice_transport_driver.F90"), ["state_to_tracers", "tracers_to_state"])
    return m, ns


def state_block(seed):
    r = np.random.default_rng(seed)
    sh = (NYB, NXB)
    aicen = r.uniform(0.02, 0.4, (NCAT,) + sh)
    aicen[r.uniform(0, 1, aicen.shape) < 0.3] = 0.0
    aicen[r.uniform(0, 1, aicen.shape) < 0.05] = 4.0e-12
    st = {"aice0": 1.0 - aicen.sum(axis=0), "aicen": aicen, "vicen": aicen * r.uniform(0.2, 3.0, aicen.shape),
          "vsnon": aicen * r.uniform(0.0, 0.4, aicen.shape), "trcrn": r.uniform(-3.0e8, 1.0, (NCAT, NTRCR) + sh)}
    return st


def run_state(m, st):
    F = fe.FArr.from_numpy
    nt = 2 + NTRCR
    aim = F(np.full((NCAT + 1, NYB, NXB), -7.0), los=[1, 1, 0])
    trm = F(np.full((NCAT, nt, NYB, NXB), -7.0))
    geo = dict(nx_block=NXB, ny_block=NYB, ntrcr=NTRCR, ntrace=nt)
    m.call("state_to_tracers", aice0=F(st["aice0"]), aicen=F(st["aicen"]), trcrn=F(st["trcrn"]), vicen=F(st["vicen"]),
           vsnon=F(st["vsnon"]), aim=aim, trm=trm, **geo)
    out = {"aim": aim.to_numpy(), "trm": trm.to_numpy()}
    # back, from a perturbed aim / trm (some areas gone to zero: those cells keep their old state)
    r = np.random.default_rng(99)
    aim2 = out["aim"] * r.uniform(0.5, 1.5, out["aim"].shape)
    aim2[r.uniform(0, 1, aim2.shape) < 0.2] = 0.0
    trm2 = out["trm"] * r.uniform(0.9, 1.1, out["trm"].shape)
    back = {k: F(st[k]) for k in ("aice0", "aicen", "trcrn", "vicen", "vsnon")}
    m.call("tracers_to_state", aim=F(aim2, los=[1, 1, 0]), trm=F(trm2), **back, **geo)
    out.update(aim2=aim2, trm2=trm2)
    out.update({f"back_{k}": v.to_numpy() for k, v in back.items()})
    return out


def main():
    m = load()
    out = {"nxb": NXB, "nyb": NYB}
    cases = [("slow_mid", 11, 0.6, 1), ("slow_first", 12, 0.6, 0), ("fast_mid", 13, 2.5, 1), ("stop", 14, 12.0, 1)]
    for name, seed, speed, midpt in cases:
        d = block(seed, speed)
        dt = 3600.0
        res = run(m, d, dt, bool(midpt))
        for k, v in d.items():
            out[f"{name}/{k}"] = v
        out[f"{name}/dt"] = dt
        out[f"{name}/l_dp_midpt"] = midpt
        for k, v in res.items():
            out[f"{name}/out/{k}"] = v
        print(name, "l_stop", res["l_stop"], "istop, jstop", res["istop"], res["jstop"],
              "max |dpx|", float(np.abs(res["dpx"]).max()))
    mf = load_fields()
    out.update(ncat=NCAT, tracer_type=np.array(TRACER_TYPE), depend=np.array(DEPEND),
               has_dependents=np.array(has_dependents(DEPEND), dtype=np.int32))
    for name, seed, open_frac in [("cf_pack", 21, 0.1), ("cf_edge", 22, 0.45)]:
        d = fields_block(seed, open_frac)
        res = run_fields(mf, d)
        for k, v in d.items():
            out[f"{name}/{k}"] = v
        for k, v in res.items():
            out[f"{name}/out/{k}"] = v
        print(name, "icells", res["icells"], "max |mx|", float(np.abs(res["mx"]).max()), "max |tx|",
              float(np.abs(res["tx"]).max()), "limited cells",
              int(np.count_nonzero(res["tx"])))
    mx3 = load_fluxes()
    for name, seed, amp, shear in [("tri_small", 31, 0.25, 0.0), ("tri_large", 32, 0.9, 0.0), ("tri_shear", 33, 0.45, 1.0)]:
        d = triangle_block(seed, amp, shear)
        for k, v in d.items():
            out[f"{name}/{k}"] = v
        for edge in ("east", "north"):
            _, tri = run_triangles(mx3, edge, d["dpx"], d["dpy"], d["dxu"], d["dyu"])
            for k, v in tri.items():
                out[f"{name}/{edge}/{k}"] = v
            print(name, edge, "cells per group", tri["icells"])
    for name, cf_seed, vel_seed, speed, open_frac in [("pipe_a", 21, 41, 1.5, 0.1), ("pipe_b", 22, 42, 3.0, 0.45)]:
        d = fields_block(cf_seed, open_frac)
        vel = block(vel_seed, speed)
        r = np.random.default_rng(vel_seed + 100)
        tarear = 1.0 / (vel["dxu"] * vel["dyu"] * r.uniform(0.95, 1.05, (NYB, NXB)))
        res = run_pipeline(mx3, d, vel, 3600.0, tarear)
        for k, v in d.items():
            out[f"{name}/{k}"] = v
        for k, v in vel.items():
            out[f"{name}/{k}"] = v
        out[f"{name}/tarear"], out[f"{name}/dt"] = tarear, 3600.0
        for k, v in res.items():
            out[f"{name}/out/{k}"] = v
        print(name, "stops", res["stops"].tolist(), "max |dmm|", float(np.abs(res["mm_new"] - d["mm"]).max()),
              "cells per group east", res["east/icells"], "north", res["north/icells"])
    md, cns = load_driver()
    out.update(ntrcr=NTRCR, nt_qsno=NT_QSNO, nslyr=NSLYR, rhos=cns["rhos"], Lfresh=cns["lfresh"])
    st = state_block(51)
    res = run_state(md, st)
    for k, v in st.items():
        out[f"state/{k}"] = v
    for k, v in res.items():
        out[f"state/out/{k}"] = v
    print("state", "cells with ice", int((res["aim"][1:] > 1e-11).sum()), "rhos*Lfresh", cns["rhos"] * cns["lfresh"])
    np.savez_compressed(OUT, **out)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
