"""ctypes wrapper around oracle/_build/libevp_oracle.so (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module (see oracle/evp_oracle.h).  PARITY PINNED TO THE REFERENCE TEXT: arithmetic, halo /
tripole code (MPI build, one task) and the whole evp driver are executed mechanically from the
reference checkout and the oracle reproduces them bit for bit (tests/test_oracle_reference_text.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libevp_oracle.so")

_DP = C.POINTER(C.c_double)
_IP = C.POINTER(C.c_int)


class Cfg(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("nx_global", "ny_global", "nslabs", "ndte", "ew_cyclic",
                                       "ns_tripole", "grid_tripole", "tilt_mode",
                                       "strength_masked", "nthreads")] + \
               [(n, C.c_double) for n in ("dt_init", "cosw", "sinw", "rhow", "rhoi", "rhos",
                                          "gravit", "a_min", "m_min")] + \
               [("revised_evp", C.c_int), ("no_halo", C.c_int)]


_GRID_R8 = ("dxt", "dyt", "dxhy", "dyhx", "cxp", "cyp", "cxm", "cym", "tarea", "uarea",
            "tarear", "uarear", "tinyarea", "fcor")


class GridS(C.Structure):
    _fields_ = [(n, _DP) for n in _GRID_R8] + [("tmask", _IP), ("umask", _IP)]


_STATE_IN = ("aice", "vice", "vsno", "aice_init", "strairxT", "strairyT", "uocn", "vocn",
             "ss_tltx", "ss_tlty", "Cw", "strength_in")
_STATE_OUT = ("strairx", "strairy", "strtltx", "strtlty", "strocnx", "strocny", "strintx",
              "strinty", "strocnxT", "strocnyT", "divu", "shear", "rdg_conv", "rdg_shear",
              "prs_sig", "uvel_init", "vvel_init", "tmass", "strength")


class StateS(C.Structure):
    _fields_ = [(n, _DP) for n in _STATE_IN] + [("uvel", _DP), ("vvel", _DP),
               ("stressp", _DP * 4), ("stressm", _DP * 4), ("stress12", _DP * 4),
               ("iceumask", _IP), ("fm", _DP)] + [(n, _DP) for n in _STATE_OUT] + \
               [("icetmask", _IP), ("strax", _DP), ("stray", _DP)]


class Stats(C.Structure):
    _fields_ = [("loop_seconds", C.c_double), ("total_seconds", C.c_double),
                ("active_t_cells", C.c_longlong), ("active_u_cells", C.c_longlong)]


def build(force: bool = False) -> str:
    if force or not os.path.exists(_SO) or \
            os.path.getmtime(_SO) < max(os.path.getmtime(os.path.join(_HERE, f))
                                        for f in ("evp_oracle.c", "strength_oracle.c", "strength_oracle.h", "remap_oracle.c", "remap_oracle.h",
                                                  "eap_oracle.inc", "eap_oracle.h", "../cice5_b200/csrc/evp_exp.h",
                                                  "../cice5_b200/csrc/evp_trig.h")):
        subprocess.run(["make", "-C", _HERE] + (["-B"] if force else []), check=True,
                       stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.evp_oracle_run.restype = C.c_int
        _lib.evp_oracle_halo_r8.restype = C.c_int
        _lib.evp_oracle_halo_i4.restype = C.c_int
    return _lib


def _dp(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_DP)


def _ip(a):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_IP)


def make_cfg(dom, ndte, dt_init, nslabs=1, nthreads=0, cosw=1.0, sinw=0.0, tilt_mode=0,
             strength_masked=0, revised_evp=False, no_halo=False) -> Cfg:
    from cice5_b200 import fixtures as fx
    return Cfg(nx_global=dom.nx, ny_global=dom.ny, nslabs=nslabs, ndte=ndte,
               ew_cyclic=int(dom.ew_cyclic), ns_tripole=int(dom.ns_tripole),
               grid_tripole=int(dom.grid_tripole), tilt_mode=tilt_mode,
               strength_masked=strength_masked, nthreads=nthreads, dt_init=dt_init,
               cosw=cosw, sinw=sinw, rhow=fx.RHOW, rhoi=fx.RHOI, rhos=fx.RHOS,
               gravit=fx.GRAVIT, a_min=1.0e-3, m_min=1.0e-2, revised_evp=int(revised_evp), no_halo=int(no_halo))


def run_case(case, nslabs=1, nthreads=0, dt_init=None, **cfg_kw):
    """One evp(dt) call on a fixtures.Case; updates case.io / case.out in place.

    Returns the Stats structure."""
    dom = case.dom
    cfg = make_cfg(dom, case.ndte, case.dt if dt_init is None else dt_init, nslabs, nthreads,
                   **cfg_kw)
    gf = case.grid.f
    g = GridS(**{n: _dp(gf[n]) for n in _GRID_R8}, tmask=_ip(gf["tmask"]), umask=_ip(gf["umask"]))
    inp, io, out = case.inp, case.io, case.out
    s = StateS()
    for n in _STATE_IN:
        setattr(s, n, _dp(inp["strength" if n == "strength_in" else n]))
    s.uvel, s.vvel, s.fm = _dp(io["uvel"]), _dp(io["vvel"]), _dp(io["fm"])
    for c in range(4):
        s.stressp[c] = _dp(io[f"stressp_{c + 1}"])
        s.stressm[c] = _dp(io[f"stressm_{c + 1}"])
        s.stress12[c] = _dp(io[f"stress12_{c + 1}"])
    s.iceumask = _ip(io["iceumask"])
    for n in _STATE_OUT:
        setattr(s, n, _dp(out["strength_out" if n == "strength" else n]))
    s.icetmask = _ip(out["icetmask"])
    if inp.get("strax") is not None:           # ACCESS build, calc_strair = .false.
        s.strax, s.stray = _dp(inp["strax"]), _dp(inp["stray"])
    st = Stats()
    rc = lib().evp_oracle_run(C.byref(cfg), C.byref(g), C.byref(s), C.c_double(case.dt), C.byref(st))
    if rc != 0:
        raise RuntimeError(f"evp_oracle_run failed with status {rc}")
    return st


def halo_r8(dom, a, loc, kind, fill=0.0, nslabs=1):
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    rc = lib().evp_oracle_halo_r8(C.byref(cfg), _dp(a), loc, kind, C.c_double(fill))
    assert rc == 0
    return a


def halo_r8_fld2(dom, a, loc, kind, fill=0.0, nslabs=1):
    """One level of the velocity pair through ice_HaloUpdate3DR8 (pair (1, nx-1) of the fold not averaged)."""
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    rc = lib().evp_oracle_halo_r8_fld2(C.byref(cfg), _dp(a), loc, kind, C.c_double(fill))
    assert rc == 0
    return a


def halo_i4(dom, a, loc, kind, fill=0, nslabs=1):
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    rc = lib().evp_oracle_halo_i4(C.byref(cfg), _ip(a), loc, kind, C.c_int(fill))
    assert rc == 0
    return a


def halo_stress(dom, a1, a2, nslabs=1):
    """ice_HaloUpdate_stress: ghosts of a1 from a2 (both global-with-ghost planes); a1 in place."""
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    rc = lib().evp_oracle_halo_stress(C.byref(cfg), _dp(a1), _dp(a2))
    assert rc == 0
    return a1


def restart_fixup(dom, io, nslabs=1):
    """ice_restart_driver.F90:370-394,431-466 on the 12 stresses of `io` (global-with-ghost planes), in place."""
    from cice5_b200 import fixtures as fx
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    arr = (_DP * 12)(*[_dp(io[n]) for n in fx.STRESS_FIELDS])
    rc = lib().evp_oracle_restart_fixup(C.byref(cfg), arr)
    assert rc == 0
    return io


def principal_stress(sp1, sm1, s121, prs):
    sig1, sig2 = np.empty_like(sp1), np.empty_like(sp1)
    lib().evp_oracle_principal_stress(C.c_longlong(sp1.size), _dp(sp1), _dp(sm1), _dp(s121),
                                      _dp(prs), _dp(sig1), _dp(sig2))
    return sig1, sig2


class StrengthCfg(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("nx", "ny", "ncat", "kstrength", "krdg_partic", "krdg_redist")] + \
               [(n, C.c_double) for n in ("mu_rdg", "Cf", "rhow", "rhoi", "gravit")] + [("exp_mode", C.c_int)]


def ice_strength(dom, aice, vice, aice0, aicen, vicen, cellmask=None, kstrength=1, krdg_partic=1,
                 krdg_redist=1, mu_rdg=3.0, Cf=17.0, cr_exp=False):
    """ice_strength (source/ice_mechred.F90:2111-2269) on global-with-ghost planes; aicen / vicen
    are (ncat, ny+2, nx+2).  cellmask (int32, non-zero = in the icellt list) or None = all cells.
    cr_exp: the correctly rounded evp_exp (what the device uses) instead of libm's exp."""
    from cice5_b200 import fixtures as fx
    ncat = aicen.shape[0]
    cfg = StrengthCfg(nx=dom.nx, ny=dom.ny, ncat=ncat, kstrength=kstrength, krdg_partic=krdg_partic,
                      krdg_redist=krdg_redist, mu_rdg=mu_rdg, Cf=Cf, rhow=fx.RHOW, rhoi=fx.RHOI,
                      gravit=fx.GRAVIT, exp_mode=1 if cr_exp else 0)
    out = np.zeros(aice.shape)
    f = lib().strength_oracle_run
    rc = f(C.byref(cfg), _dp(aice), _dp(vice), _dp(aice0), _dp(np.ascontiguousarray(aicen)),
           _dp(np.ascontiguousarray(vicen)), _ip(cellmask) if cellmask is not None else None, _dp(out))
    if rc != 0:
        raise RuntimeError(f"strength_oracle_run failed with status {rc}")
    return out


def cr_exp(x):
    """evp_exp (cice5_b200/csrc/evp_exp.h, the exp shared by the device kernels and this oracle) on an array."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty_like(x)
    lib().strength_oracle_cr_exp(C.c_longlong(x.size), _dp(x), _dp(y))
    return y


def construct_fields(dom, tracer_type, depend, has_dependents, hm, xav, yav, xxav, yyav, mm, tm, no_halo=False, nslabs=1,
                     masks=False):
    """make_masks + construct_fields + limited_gradient (source/ice_transport_remap.F90:867-1484) for categories
    0..ncat, then (unless no_halo) the ghost updates of horizontal_remap (:579-584, :616-621).  mm: (ncat+1, ny+2, nx+2);
    tm: (ncat, ntrace, ny+2, nx+2).  Returns a dict mc, mx, my, tc, tx, ty (+ mmask, tmask, icells with masks=True)."""
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    mm, tm = np.ascontiguousarray(mm, dtype=np.float64), np.ascontiguousarray(tm, dtype=np.float64)
    ncat, ntrace = tm.shape[0], tm.shape[1]
    assert mm.shape[0] == ncat + 1
    ia = lambda v: np.ascontiguousarray(v, dtype=np.int32)
    tt, dp, hd = ia(tracer_type), ia(depend), ia(has_dependents)
    o = {k: np.full(mm.shape, np.nan) for k in ("mc", "mx", "my")}
    o.update({k: np.full(tm.shape, np.nan) for k in ("tc", "tx", "ty")})
    if masks:
        o["mmask"], o["tmask"] = np.full(mm.shape, np.nan), np.full(tm.shape, np.nan)
        o["icells"] = np.zeros(ncat + 1, dtype=np.int32)
    rc = lib().remap_oracle_construct_fields(
        C.byref(cfg), C.c_int(ncat), C.c_int(ntrace), _ip(tt), _ip(dp), _ip(hd), C.c_int(int(no_halo)), _dp(hm), _dp(xav),
        _dp(yav), _dp(xxav), _dp(yyav), _dp(mm), _dp(tm), _dp(o["mc"]), _dp(o["mx"]), _dp(o["my"]), _dp(o["tc"]),
        _dp(o["tx"]), _dp(o["ty"]), _dp(o["mmask"]) if masks else None, _dp(o["tmask"]) if masks else None,
        _ip(o["icells"]) if masks else None)
    if rc != 0:
        raise RuntimeError(f"remap_oracle_construct_fields failed with status {rc}")
    return o


def triangles(north, dpx, dpy, dxu, dyu, coords=True):
    """locate_triangles (+ triangle_coordinates, integral_order 3) of source/ice_transport_remap.F90:1680-3187 for one
    edge orientation on one block (ny_block, nx_block).  Returns a dict xp, yp (6, 4, ny, nx), iflux, jflux (Fortran
    indices), triarea (6, ny, nx), icells (6), edgearea."""
    nyb, nxb = dpx.shape
    o = {"xp": np.full((6, 4, nyb, nxb), np.nan), "yp": np.full((6, 4, nyb, nxb), np.nan),
         "iflux": np.zeros((6, nyb, nxb), dtype=np.int32), "jflux": np.zeros((6, nyb, nxb), dtype=np.int32),
         "triarea": np.full((6, nyb, nxb), np.nan), "icells": np.zeros(6, dtype=np.int32), "edgearea": np.zeros((nyb, nxb))}
    rc = lib().remap_oracle_triangles(C.c_int(nxb), C.c_int(nyb), C.c_int(int(north)), C.c_int(int(coords)), _dp(dpx), _dp(dpy),
                                      _dp(dxu), _dp(dyu), _dp(o["xp"]), _dp(o["yp"]), _ip(o["iflux"]), _ip(o["jflux"]),
                                      _dp(o["triarea"]), _ip(o["icells"]), _dp(o["edgearea"]))
    if rc != 0:
        raise RuntimeError(f"remap_oracle_triangles failed with status {rc}")
    return o


def transport_integrals(tracer_type, depend, tri, mc, mx, my, tc=None, tx=None, ty=None):
    """transport_integrals (:3199-3509, integral_order 3) of one category from the arrays of :func:`triangles`.
    Returns mflx (ny, nx) and, with tracers, mtflx (ntrace, ny, nx)."""
    nyb, nxb = mc.shape
    ia = lambda v: np.ascontiguousarray(v, dtype=np.int32)
    tt, dp = ia(tracer_type), ia(depend)
    mflx = np.full((nyb, nxb), np.nan)
    mtflx = np.full(tc.shape, np.nan) if tc is not None else None
    rc = lib().remap_oracle_transport_integrals(
        C.c_int(nxb), C.c_int(nyb), C.c_int(tt.size), _ip(tt), _ip(dp), _dp(tri["xp"]), _dp(tri["yp"]), _ip(tri["iflux"]),
        _ip(tri["jflux"]), _dp(tri["triarea"]), _dp(mc), _dp(mx), _dp(my), _dp(mflx),
        _dp(np.ascontiguousarray(tc)) if tc is not None else None, _dp(np.ascontiguousarray(tx)) if tc is not None else None,
        _dp(np.ascontiguousarray(ty)) if tc is not None else None, _dp(mtflx) if tc is not None else None)
    if rc != 0:
        raise RuntimeError(f"remap_oracle_transport_integrals failed with status {rc}")
    return mflx, mtflx


def update_fields(tracer_type, depend, tarear, mflxe, mflxn, mm, mtflxe=None, mtflxn=None, tm=None):
    """update_fields (:3517-3729) of one category; mm (and tm) are updated in place.  Returns (l_stop, istop, jstop)."""
    nyb, nxb = mm.shape
    ia = lambda v: np.ascontiguousarray(v, dtype=np.int32)
    tt, dp = ia(tracer_type), ia(depend)
    ls, i_s, j_s = C.c_int(0), C.c_int(0), C.c_int(0)
    rc = lib().remap_oracle_update_fields(C.c_int(nxb), C.c_int(nyb), C.c_int(tt.size), _ip(tt), _ip(dp), _dp(tarear), _dp(mflxe),
                                          _dp(mflxn), _dp(mm), _dp(mtflxe) if tm is not None else None,
                                          _dp(mtflxn) if tm is not None else None, _dp(tm) if tm is not None else None,
                                          C.byref(ls), C.byref(i_s), C.byref(j_s))
    if rc != 0:
        raise RuntimeError(f"remap_oracle_update_fields failed with status {rc}")
    return bool(ls.value), i_s.value, j_s.value


def horizontal_remap(dom, dt, tracer_type, depend, has_dependents, g, uvel, vvel, mm, tm, no_halo=False, nslabs=1):
    """All of horizontal_remap (source/ice_transport_remap.F90:309-850; l_fixed_area = F, integral_order = 3, l_dp_midpt = T)
    on global-with-ghost planes.  g: dict hm, xav, yav, xxav, yyav, dxu, dyu, HTN, HTE, tarear.  mm (ncat+1, ny+2, nx+2)
    and tm (ncat, ntrace, ny+2, nx+2) are updated in place on the physical cells.  Returns (stop, category, istop, jstop):
    stop 0 = completed, 1 = bad departure points, 2 / 3 = negative open-water / ice area (the reference aborts)."""
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    ncat, ntrace = tm.shape[0], tm.shape[1]
    ia = lambda v: np.ascontiguousarray(v, dtype=np.int32)
    tt, dp, hd = ia(tracer_type), ia(depend), ia(has_dependents)
    st, sc, i_s, j_s = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
    rc = lib().remap_oracle_horizontal_remap(
        C.byref(cfg), C.c_double(dt), C.c_int(ncat), C.c_int(ntrace), _ip(tt), _ip(dp), _ip(hd), C.c_int(int(no_halo)),
        *[_dp(np.ascontiguousarray(g[k], dtype=np.float64)) for k in ("hm", "xav", "yav", "xxav", "yyav", "dxu", "dyu", "HTN",
                                                                      "HTE", "tarear")],
        _dp(uvel), _dp(vvel), _dp(mm), _dp(tm), C.byref(st), C.byref(sc), C.byref(i_s), C.byref(j_s))
    if rc != 0:
        raise RuntimeError(f"remap_oracle_horizontal_remap failed with status {rc}")
    return st.value, sc.value, i_s.value, j_s.value


def state_to_tracers(st, nt_qsno, nslyr, rhos, Lfresh):
    """state_to_tracers (source/ice_transport_driver.F90:789-900) of one block.  st: dict aice0 (ny, nx), aicen / vicen /
    vsnon (ncat, ny, nx), trcrn (ncat, ntrcr, ny, nx).  Returns aim (ncat+1, ny, nx), trm (ncat, 2+ntrcr, ny, nx)."""
    a = {k: np.ascontiguousarray(st[k], dtype=np.float64) for k in ("aice0", "aicen", "trcrn", "vicen", "vsnon")}
    ncat, ntrcr = a["trcrn"].shape[0], a["trcrn"].shape[1]
    nyb, nxb = a["aice0"].shape
    aim, trm = np.full((ncat + 1, nyb, nxb), np.nan), np.full((ncat, 2 + ntrcr, nyb, nxb), np.nan)
    rc = lib().remap_oracle_state_to_tracers(C.c_int(nxb), C.c_int(nyb), C.c_int(ncat), C.c_int(ntrcr), C.c_int(nt_qsno),
                                             C.c_int(nslyr), C.c_double(rhos), C.c_double(Lfresh), _dp(a["aice0"]), _dp(a["aicen"]),
                                             _dp(a["trcrn"]), _dp(a["vicen"]), _dp(a["vsnon"]), _dp(aim), _dp(trm))
    assert rc == 0
    return aim, trm


def tracers_to_state(aim, trm, st, nt_qsno, nslyr, rhos, Lfresh):
    """tracers_to_state (:908-1002): the state dict `st` (as for state_to_tracers, C-contiguous float64) is updated in place."""
    ncat, ntrcr = st["trcrn"].shape[0], st["trcrn"].shape[1]
    nyb, nxb = st["aice0"].shape
    rc = lib().remap_oracle_tracers_to_state(C.c_int(nxb), C.c_int(nyb), C.c_int(ncat), C.c_int(ntrcr), C.c_int(nt_qsno),
                                             C.c_int(nslyr), C.c_double(rhos), C.c_double(Lfresh), _dp(aim), _dp(trm),
                                             _dp(st["aice0"]), _dp(st["aicen"]), _dp(st["trcrn"]), _dp(st["vicen"]), _dp(st["vsnon"]))
    assert rc == 0
    return st


def transport_remap(dom, dt, tracer_type, depend, has_dependents, g, uvel, vvel, st, nt_qsno, nslyr, rhos, Lfresh,
                    no_halo=False, nslabs=1):
    """transport_remap (source/ice_transport_driver.F90:198-627, checks off) on global-with-ghost planes: the state dict `st`
    (aice0, aicen, trcrn, vicen, vsnon; C-contiguous float64) is updated in place, ghost cells by bound_state.  Returns
    (stop, category, istop, jstop) as :func:`horizontal_remap`."""
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    ncat, ntrcr = st["trcrn"].shape[0], st["trcrn"].shape[1]
    ia = lambda v: np.ascontiguousarray(v, dtype=np.int32)
    tt, dp, hd = ia(tracer_type), ia(depend), ia(has_dependents)
    assert tt.size == 2 + ntrcr
    sv, sc, i_s, j_s = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
    rc = lib().remap_oracle_transport_remap(
        C.byref(cfg), C.c_double(dt), C.c_int(ncat), C.c_int(ntrcr), C.c_int(nt_qsno), C.c_int(nslyr), C.c_double(rhos),
        C.c_double(Lfresh), _ip(tt), _ip(dp), _ip(hd), C.c_int(int(no_halo)),
        *[_dp(np.ascontiguousarray(g[k], dtype=np.float64)) for k in ("hm", "xav", "yav", "xxav", "yyav", "dxu", "dyu", "HTN",
                                                                      "HTE", "tarear")],
        _dp(uvel), _dp(vvel), _dp(st["aice0"]), _dp(st["aicen"]), _dp(st["trcrn"]), _dp(st["vicen"]), _dp(st["vsnon"]),
        C.byref(sv), C.byref(sc), C.byref(i_s), C.byref(j_s))
    if rc != 0:
        raise RuntimeError(f"remap_oracle_transport_remap failed with status {rc}")
    return sv.value, sc.value, i_s.value, j_s.value


def departure_points(dom, dt, uvel, vvel, dxu, dyu, HTN, HTE, l_dp_midpt=True, no_halo=False, nslabs=1):
    """departure_points (source/ice_transport_remap.F90:1493-1670) + the ghost update of dpx, dpy that
    horizontal_remap issues (:574-578), on global-with-ghost planes.  Returns (dpx, dpy, l_stop, istop, jstop)."""
    cfg = make_cfg(dom, 1, 1.0, nslabs)
    dpx, dpy = np.empty_like(uvel), np.empty_like(uvel)
    ls, i_s, j_s = C.c_int(0), C.c_int(0), C.c_int(0)
    rc = lib().remap_oracle_departure_points(C.byref(cfg), C.c_double(dt), C.c_int(int(l_dp_midpt)), C.c_int(int(no_halo)),
                                             _dp(uvel), _dp(vvel), _dp(dxu), _dp(dyu), _dp(HTN), _dp(HTE), _dp(dpx), _dp(dpy),
                                             C.byref(ls), C.byref(i_s), C.byref(j_s))
    if rc != 0:
        raise RuntimeError(f"remap_oracle_departure_points failed with status {rc}")
    return dpx, dpy, bool(ls.value), i_s.value, j_s.value


# ---- EAP (kdyn = 2): oracle/eap_oracle.h ------------------------------------------------------------------------
EAP_TABLE_SHAPE = (21, 41, 41)          # (na_yield, ny_yield, nx_yield): C order of Fortran (nx_yield, ny_yield, na_yield)
EAP_TABLES = ("s11r", "s12r", "s22r", "s11s", "s12s", "s22s")
EAP_CORNER = tuple(f"a11_{c}" for c in (1, 2, 3, 4)) + tuple(f"a12_{c}" for c in (1, 2, 3, 4))
EAP_AVG = ("a11", "a12")
EAP_OUT = ("e11", "e12", "e22", "s11", "s12", "s22", "yieldstress11", "yieldstress12", "yieldstress22")


class EapS(C.Structure):
    _fields_ = [(n, _DP) for n in EAP_TABLES] + [("a11_c", _DP * 4), ("a12_c", _DP * 4)] + \
               [(n, _DP) for n in EAP_AVG + EAP_OUT] + [("trig_mode", C.c_int)]


_eap_tables = None


def eap_tables(ia=(1, 21), ix=(1, 41), iy=(1, 41)):
    """init_eap's lookup tables (source/ice_dyn_eap.F90:557-618) as a dict of (21, 41, 41) arrays; the full tables
    are computed once per process."""
    global _eap_tables
    full = ia == (1, 21) and ix == (1, 41) and iy == (1, 41)
    if full and _eap_tables is not None:
        return _eap_tables
    t = {n: np.zeros(EAP_TABLE_SHAPE) for n in EAP_TABLES}
    f = lib().eap_oracle_init_tables
    f.restype = None
    f(*[_dp(t[n]) for n in EAP_TABLES], *[C.c_int(v) for v in (*ia, *ix, *iy)])
    if full:
        _eap_tables = t
    return t


def eap_state(dom):
    """Fresh EAP state as init_eap leaves it (:531-548): a11_k = 0.5, a12_k = 0, everything else 0."""
    sh = (dom.ny + 2, dom.nx + 2)
    e = {n: np.zeros(sh) for n in EAP_CORNER + EAP_AVG + EAP_OUT}
    for c in (1, 2, 3, 4):
        e[f"a11_{c}"][...] = 0.5
    return e


def _eap_struct(tables, eap, trig_mode):
    E = EapS(trig_mode=trig_mode)
    for n in EAP_TABLES:
        setattr(E, n, _dp(tables[n]))
    for c in range(4):
        E.a11_c[c] = _dp(eap[f"a11_{c + 1}"])
        E.a12_c[c] = _dp(eap[f"a12_{c + 1}"])
    for n in EAP_AVG + EAP_OUT:
        setattr(E, n, _dp(eap[n]))
    return E


def run_case_eap(case, eap, tables=None, nslabs=1, nthreads=0, dt_init=None, trig_mode=1, **cfg_kw):
    """One eap(dt) call on a fixtures.Case + an EAP state dict (eap_state); updates case.io / case.out / eap in place."""
    dom = case.dom
    tables = eap_tables() if tables is None else tables
    cfg = make_cfg(dom, case.ndte, case.dt if dt_init is None else dt_init, nslabs, nthreads, **cfg_kw)
    gf = case.grid.f
    g = GridS(**{n: _dp(gf[n]) for n in _GRID_R8}, tmask=_ip(gf["tmask"]), umask=_ip(gf["umask"]))
    inp, io, out = case.inp, case.io, case.out
    s = StateS()
    for n in _STATE_IN:
        setattr(s, n, _dp(inp["strength" if n == "strength_in" else n]))
    s.uvel, s.vvel, s.fm = _dp(io["uvel"]), _dp(io["vvel"]), _dp(io["fm"])
    for c in range(4):
        s.stressp[c] = _dp(io[f"stressp_{c + 1}"])
        s.stressm[c] = _dp(io[f"stressm_{c + 1}"])
        s.stress12[c] = _dp(io[f"stress12_{c + 1}"])
    s.iceumask = _ip(io["iceumask"])
    for n in _STATE_OUT:
        setattr(s, n, _dp(out["strength_out" if n == "strength" else n]))
    s.icetmask = _ip(out["icetmask"])
    E = _eap_struct(tables, eap, trig_mode)
    st = Stats()
    f = lib().eap_oracle_run
    f.restype = C.c_int
    rc = f(C.byref(cfg), C.byref(g), C.byref(s), C.byref(E), C.c_double(case.dt), C.byref(st))
    if rc != 0:
        raise RuntimeError(f"eap_oracle_run failed with status {rc}")
    return st


def trig(name):
    """evp_trig.h's sin / cos / atan2 as Python callables."""
    f = getattr(lib(), f"evp_oracle_{name}")
    f.restype = C.c_double
    f.argtypes = [C.c_double] * (2 if name == "atan2" else 1)
    return f
