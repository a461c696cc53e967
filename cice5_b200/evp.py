"""Host-side mirror of the reference interface for the EVP path, over the C ABI (ctypes).

The reference's entry point is ``call evp(dt)`` in ``module ice_dyn_evp``
(``source/ice_dyn_evp.F90:68``) operating on module arrays; :class:`EvpGpu` holds the same
arrays by their reference names (``uvel``, ``stressp_1`` ...) and :meth:`EvpGpu.evp` is that
call.  Everything numeric happens in ``lib/libevp_b200.so`` (CUDA, sm_100a); if the library or a
GPU is missing this module raises -- there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

_DP = C.POINTER(C.c_double)
_IP = C.POINTER(C.c_int)


class EvpError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "nx_global", "ny_global", "nx_block", "ny_block", "rank", "nranks", "device", "ndte",
        "kdyn", "revised_evp", "ew_boundary", "ns_boundary", "grid_tripole", "tilt_mode",
        "strength_masked", "use_cuda_graph")] + [(n, C.c_double) for n in (
        "dt_init", "cosw", "sinw", "rhow", "rhoi", "rhos", "gravit", "a_min", "m_min", "xmin")] + [
        ("slab_start", C.c_int * 9)]


GRID_R8 = ("dxt", "dyt", "dxhy", "dyhx", "cxp", "cyp", "cxm", "cym", "tarea", "uarea", "tarear",
           "uarear", "tinyarea", "fcor")
IN_R8 = ("aice", "vice", "vsno", "aice_init", "strairxT", "strairyT", "uocn", "vocn", "ss_tltx",
         "ss_tlty", "Cdn_ocn", "strength")
OUT_R8 = ("strairx", "strairy", "strtltx", "strtlty", "strocnx", "strocny", "strintx", "strinty",
          "strocnxT", "strocnyT", "divu", "shear", "rdg_conv", "rdg_shear", "prs_sig",
          "uvel_init", "vvel_init", "tmass", "strength")


class GridS(C.Structure):
    _fields_ = [(n, _DP) for n in GRID_R8] + [("tmask", _IP), ("umask", _IP), ("HTE", _DP), ("HTN", _DP)]


class ItdS(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("ncat", "kstrength", "krdg_partic", "krdg_redist")] + \
               [("mu_rdg", C.c_double), ("Cf", C.c_double), ("aicen", _DP), ("vicen", _DP), ("aice0", _DP)]


class InS(C.Structure):
    _fields_ = [(n, _DP) for n in IN_R8] + [("strax", _DP), ("stray", _DP)]


class InOutS(C.Structure):
    _fields_ = [("uvel", _DP), ("vvel", _DP), ("stressp", _DP * 4), ("stressm", _DP * 4),
                ("stress12", _DP * 4), ("iceumask", _IP), ("fm", _DP)]


class OutS(C.Structure):
    _fields_ = [(n, _DP) for n in OUT_R8] + [("icetmask", _IP)]


# EAP (kdyn = 2): the module variables of source/ice_dyn_eap.F90:31-57
EAP_TABLES = ("s11r", "s12r", "s22r", "s11s", "s12s", "s22s")
EAP_TABLE_SHAPE = (21, 41, 41)      # C order of Fortran (nx_yield, ny_yield, na_yield)
EAP_CORNER = tuple(f"a11_{c}" for c in (1, 2, 3, 4)) + tuple(f"a12_{c}" for c in (1, 2, 3, 4))
EAP_AVG = ("a11", "a12")
EAP_OUT = ("e11", "e12", "e22", "s11", "s12", "s22", "yieldstress11", "yieldstress12", "yieldstress22")


class EapTablesS(C.Structure):
    _fields_ = [(n, _DP) for n in EAP_TABLES]


class EapStateS(C.Structure):
    _fields_ = [("a11", _DP * 4), ("a12", _DP * 4), ("a11_avg", _DP), ("a12_avg", _DP)]


class EapOutS(C.Structure):
    _fields_ = [(n, _DP) for n in EAP_OUT]


class Stats(C.Structure):
    _fields_ = [("prep_ms", C.c_double), ("loop_ms", C.c_double), ("finish_ms", C.c_double),
                ("total_ms", C.c_double), ("active_t_cells", C.c_longlong),
                ("active_u_cells", C.c_longlong), ("kernel_launches", C.c_longlong),
                ("loop_kernel_launches", C.c_longlong)]


EXPORTS = ("evp_gpu_init", "evp_gpu_set_grid", "evp_gpu_run", "evp_gpu_upload", "evp_gpu_compute",
           "evp_gpu_download", "evp_gpu_compute_group", "evp_gpu_get_stats", "evp_gpu_synchronize", "evp_gpu_peer_blob_size",
           "evp_gpu_get_peer_blob", "evp_gpu_set_peers", "evp_gpu_finalize", "evp_gpu_strerror",
           "evp_gpu_last_cuda_error", "evp_gpu_device_count", "evp_gpu_get_trace", "evp_gpu_upload_itd",
           "evp_gpu_principal_stress", "evp_gpu_pin_host_arrays", "evp_gpu_set_resident",
           "evp_gpu_invalidate_state", "evp_gpu_restart_fixup", "evp_gpu_restart_fixup_group",
           "evp_gpu_principal_stress_static", "evp_gpu_set_transport_grid", "evp_gpu_departure_points_compute",
           "evp_gpu_departure_points_group", "evp_gpu_departure_points_fetch", "evp_gpu_departure_points",
           "evp_gpu_set_transport_tracers", "evp_gpu_construct_fields_upload", "evp_gpu_construct_fields_compute",
           "evp_gpu_construct_fields_group", "evp_gpu_construct_fields_fetch", "evp_gpu_construct_fields",
           "evp_gpu_get_transport_timing", "evp_gpu_horizontal_remap_compute", "evp_gpu_horizontal_remap_group",
           "evp_gpu_horizontal_remap_fetch", "evp_gpu_horizontal_remap", "evp_gpu_transport_remap_upload",
           "evp_gpu_transport_remap_compute", "evp_gpu_transport_remap_group", "evp_gpu_transport_remap_fetch",
           "evp_gpu_transport_remap", "evp_gpu_set_eap_tables", "evp_gpu_upload_eap", "evp_gpu_download_eap")
REMAP_MAX_TRACE, REMAP_MAX_PARENTS = 128, 16     # EVP_REMAP_MAX_TRACE, EVP_REMAP_MAX_PARENTS of include/evp_b200.h
STATIC_MAX_CELLS = 131072       # EVP_STATIC_MAX_CELLS of include/evp_b200.h

# download groups of the resident-state protocol (include/evp_b200.h)
DL_VELOCITY, DL_STRESS, DL_DEFORM, DL_OCN_T, DL_HISTORY, DL_ALL = 1, 2, 4, 8, 16, 31

_libs: dict = {}


def load_library(strict: bool = True):
    """dlopen the in-tree CUDA library (building it if nvcc is here and it is stale).

    strict=True (the product default) is the build without FMA contraction: it evaluates the
    reference's IEEE operations one by one and is bitwise equal to the CPU oracle.  strict=False
    is the same source with contraction on, kept for comparison only (EVP at ndte=120 amplifies
    round-off on some grids, so a contracted build cannot promise 1e-10 there)."""
    if strict in _libs:
        return _libs[strict]
    path = os.environ.get("EVP_B200_LIB") or _build.lib_path(strict)   # override: kernel-tuning experiments
    if not os.environ.get("EVP_B200_LIB"):
        # refuse a stale library: a compile failure of edited sources must not let the tests and the bench run
        # (and report parity for) code that no longer builds; only "no nvcc on this box" is tolerated, and only
        # when the library on disk is newer than every source (the GPU box runs the prebuilt file)
        try:
            _build.build()
        except _build.NoCompiler:
            if not os.path.exists(path) or _build.is_stale(path):
                raise
    if not os.path.exists(path):
        raise EvpError(f"CUDA extension {path} is missing: build it with `python -m cice5_b200.build`")
    lib = C.CDLL(path)
    lib.evp_gpu_strerror.restype = C.c_char_p
    lib.evp_gpu_last_cuda_error.restype = C.c_char_p
    for n in EXPORTS:
        getattr(lib, n)
    _libs[strict] = lib
    return lib


def _dp(a):
    if a is None:
        return None
    if a.dtype != np.float64 or not a.flags["C_CONTIGUOUS"]:
        raise EvpError("arrays must be C-contiguous float64 with shape (ny_block, nx_block)")
    return a.ctypes.data_as(_DP)


def _ip(a):
    if a is None:
        return None
    if a.dtype != np.int32 or not a.flags["C_CONTIGUOUS"]:
        raise EvpError("mask arrays must be C-contiguous int32")
    return a.ctypes.data_as(_IP)


class EvpGpu:
    """One task's EVP solver on one GPU (one block, x-slab ``rank`` of ``nranks``)."""

    def __init__(self, nx_global, ny_global, ndte, dt_init, *, rank=0, nranks=1, device=0,
                 ew_cyclic=True, ns_boundary="open", grid_tripole=False, tilt_mode=0,
                 strength_masked=0, cosw=1.0, sinw=0.0, rhow=1026.0, rhoi=917.0, rhos=330.0,
                 gravit=9.80616, a_min=1.0e-3, m_min=1.0e-2, kdyn=1, revised_evp=False,
                 use_cuda_graph=False, strict=True, compact_metrics=True, slab_start=None, xmin=0.0):
        self.lib = load_library(strict)
        self.compact_metrics = compact_metrics      # pass HTE / HTN to set_grid when the grid has them
        ns = {"open": 0, "closed": 1, "tripole": 2}[ns_boundary]
        if slab_start is None:
            if nx_global % nranks:
                raise EvpError("nranks must divide nx_global (or pass slab_start)")
            slab_start = [r * (nx_global // nranks) for r in range(nranks)]
        slab_start = [int(x) for x in slab_start]
        if len(slab_start) != nranks or slab_start[0] != 0:
            raise EvpError("slab_start needs one 0-based first column per rank, starting at 0")
        self.slab_start = slab_start
        self.nxl = (slab_start + [nx_global])[rank + 1] - slab_start[rank]
        self.shape = (ny_global + 2, self.nxl + 2)
        self.cfg = Config(nx_global=nx_global, ny_global=ny_global, nx_block=self.nxl + 2,
                          ny_block=ny_global + 2, rank=rank, nranks=nranks, device=device,
                          ndte=ndte, kdyn=kdyn, revised_evp=int(revised_evp),
                          ew_boundary=int(ew_cyclic), ns_boundary=ns,
                          grid_tripole=int(grid_tripole), tilt_mode=tilt_mode,
                          strength_masked=strength_masked, use_cuda_graph=int(use_cuda_graph),
                          dt_init=dt_init, cosw=cosw, sinw=sinw, rhow=rhow, rhoi=rhoi, rhos=rhos,
                          gravit=gravit, a_min=a_min, m_min=m_min, xmin=xmin)
        for r, x in enumerate(slab_start):
            self.cfg.slab_start[r] = x
        self._h = C.c_void_p()
        self._check(self.lib.evp_gpu_init(C.byref(self.cfg), C.byref(self._h)))
        self._keep = []

    # -- error behaviour of the reference: abort_ice(message); here: raise with the same text
    def _check(self, rc):
        if rc != 0:
            msg = self.lib.evp_gpu_strerror(rc).decode()
            if rc == 3:
                msg += " [" + self.lib.evp_gpu_last_cuda_error().decode() + "]"
            raise EvpError(msg)

    @classmethod
    def for_domain(cls, dom, ndte, dt_init, **kw):
        ns = "tripole" if dom.ns_tripole else ("open" if dom.ns_open else "closed")
        return cls(dom.nx, dom.ny, ndte, dt_init, ew_cyclic=dom.ew_cyclic, ns_boundary=ns,
                   grid_tripole=dom.grid_tripole, **kw)

    def set_grid(self, f: dict):
        for n in GRID_R8 + ("tmask", "umask"):
            if f[n].shape != self.shape:
                raise EvpError(f"grid array {n} has shape {f[n].shape}, expected {self.shape}")
        g = GridS(**{n: _dp(f[n]) for n in GRID_R8}, tmask=_ip(f["tmask"]), umask=_ip(f["umask"]))
        if f.get("HTE") is not None and f.get("HTN") is not None and self.compact_metrics:
            g.HTE, g.HTN = _dp(f["HTE"]), _dp(f["HTN"])     # optional: compact metric rows
        self._check(self.lib.evp_gpu_set_grid(self._h, C.byref(g)))

    @staticmethod
    def _in_struct(inp):
        s = InS()
        for n in IN_R8:
            key = "Cw" if n == "Cdn_ocn" and "Cdn_ocn" not in inp else n
            setattr(s, n, _dp(inp.get(key)) if n == "strength" else _dp(inp[key]))   # strength: optional
        if inp.get("strax") is not None or inp.get("stray") is not None:
            # ACCESS build with calc_strair = .false. (source/ice_dyn_evp.F90:220-229)
            s.strax, s.stray = _dp(inp.get("strax")), _dp(inp.get("stray"))
        return s

    @staticmethod
    def _io_struct(io):
        s = InOutS()
        s.uvel, s.vvel, s.fm = _dp(io["uvel"]), _dp(io["vvel"]), _dp(io["fm"])
        for c in range(4):
            s.stressp[c] = _dp(io[f"stressp_{c + 1}"])
            s.stressm[c] = _dp(io[f"stressm_{c + 1}"])
            s.stress12[c] = _dp(io[f"stress12_{c + 1}"])
        s.iceumask = _ip(io["iceumask"])
        return s

    @staticmethod
    def _out_struct(out):
        s = OutS()
        for n in OUT_R8:
            key = "strength_out" if n == "strength" else n
            if out.get(key) is not None:
                setattr(s, n, _dp(out[key]))
        if out.get("icetmask") is not None:
            s.icetmask = _ip(out["icetmask"])
        return s

    def evp(self, dt, inp, io, out):
        """``call evp(dt)``: host arrays in, host arrays out (H2D + device solve + D2H)."""
        i, s, o = self._in_struct(inp), self._io_struct(io), self._out_struct(out)
        self._check(self.lib.evp_gpu_run(self._h, C.c_double(dt), C.byref(i), C.byref(s), C.byref(o)))

    def pin_host(self, inp=None, io=None, out=None):
        """Page-lock the caller's arrays once (evp_gpu_pin_host_arrays): for hosts whose arrays keep their
        addresses, as the reference's static module arrays do.  Already page-locked arrays are left alone."""
        si = self._in_struct_partial(inp) if inp is not None else None
        sio = self._io_struct(io) if io is not None else None
        so = self._out_struct(out) if out is not None else None
        self.__dict__.setdefault("_keep_pinned", []).append((inp, io, out))     # the arrays must outlive the handle
        self._check(self.lib.evp_gpu_pin_host_arrays(self._h, C.byref(si) if si else None,
                                                     C.byref(sio) if sio else None, C.byref(so) if so else None))

    @staticmethod
    def _in_struct_partial(inp):
        s = InS()
        for n in IN_R8 + ("strax", "stray"):
            key = "Cw" if n == "Cdn_ocn" and "Cdn_ocn" not in inp else n
            setattr(s, n, _dp(inp.get(key)))
        return s

    def upload_itd(self, aicen, vicen, aice0, kstrength=1, krdg_partic=1, krdg_redist=1, mu_rdg=3.0,
                   Cf=17.0):
        """Thickness distribution for the device ice_strength (source/ice_mechred.F90:2111-2269; the
        call evp makes at source/ice_dyn_evp.F90:291-301).  aicen / vicen: (ncat, ny_block, nx_block).
        Afterwards `strength` may be missing (None) from the per-call inputs."""
        aicen = np.ascontiguousarray(aicen)
        vicen = np.ascontiguousarray(vicen)
        if aicen.shape != vicen.shape or aicen.shape[1:] != self.shape or aice0.shape != self.shape:
            raise EvpError(f"ITD arrays must be (ncat, {self.shape[0]}, {self.shape[1]})")
        s = ItdS(ncat=aicen.shape[0], kstrength=kstrength, krdg_partic=krdg_partic,
                 krdg_redist=krdg_redist, mu_rdg=mu_rdg, Cf=Cf, aicen=_dp(aicen), vicen=_dp(vicen),
                 aice0=_dp(aice0))
        self._check(self.lib.evp_gpu_upload_itd(self._h, C.byref(s)))
        self._keep = [aicen, vicen, aice0]           # copies are asynchronous on the handle's stream

    def upload(self, inp, io):
        i, s = self._in_struct(inp), self._io_struct(io)
        self._check(self.lib.evp_gpu_upload(self._h, C.byref(i), C.byref(s)))

    def compute(self, dt):
        self._check(self.lib.evp_gpu_compute(self._h, C.c_double(dt)))

    def download(self, io, out=None):
        """Resident state -> host arrays; io or out may be None (nothing of that group is copied)."""
        s = self._io_struct(io) if io is not None else None
        o = self._out_struct(out) if out is not None else None
        self._check(self.lib.evp_gpu_download(self._h, C.byref(s) if s is not None else None,
                                              C.byref(o) if o is not None else None))

    def set_resident(self, resident=True, download_mask=DL_VELOCITY | DL_DEFORM | DL_OCN_T):
        """Resident-state protocol of evp(): the dynamic state goes up only when the device does not hold it;
        only the groups of `download_mask` come back (default: what transport, ridging and the coupler read)."""
        self._check(self.lib.evp_gpu_set_resident(self._h, C.c_int(int(resident)), C.c_uint(download_mask)))

    def invalidate_state(self):
        self._check(self.lib.evp_gpu_invalidate_state(self._h))

    def restart_fixup(self):
        """What restartfile does to freshly read stresses (ice_restart_driver.F90:370-394,431-466), in place."""
        self._check(self.lib.evp_gpu_restart_fixup(self._h))

    def set_transport_grid(self, f):
        """dxu, dyu, HTN, HTE of this block (evp_gpu_set_transport_grid): needed once before departure_points."""
        g = (_DP * 4)(_dp(f["dxu"]), _dp(f["dyu"]), _dp(f["HTN"]), _dp(f["HTE"]))
        self._keep.append([f[k] for k in ("dxu", "dyu", "HTN", "HTE")])
        self._check(self.lib.evp_gpu_set_transport_grid(self._h, g))

    def departure_points_fetch(self):
        dpx, dpy = np.zeros(self.shape), np.zeros(self.shape)
        ls, i_s, j_s = C.c_int(0), C.c_int(0), C.c_int(0)
        self._check(self.lib.evp_gpu_departure_points_fetch(self._h, _dp(dpx), _dp(dpy), C.byref(ls), C.byref(i_s),
                                                            C.byref(j_s)))
        return dpx, dpy, bool(ls.value), i_s.value, j_s.value

    def departure_points(self, dt, l_dp_midpt=True):
        """departure_points (source/ice_transport_remap.F90:1493-1670) + the ghost update of dpx, dpy, from the resident
        uvel, vvel.  Returns (dpx, dpy, l_stop, istop, jstop)."""
        self._check(self.lib.evp_gpu_departure_points_compute(self._h, C.c_double(dt), C.c_int(int(l_dp_midpt))))
        return self.departure_points_fetch()

    def set_transport_tracers(self, ncat, tracer_type, depend, has_dependents, f):
        """init_transport's tracer tables (source/ice_transport_driver.F90:82-118) and the block's hm, xav, yav, xxav, yyav
        (evp_gpu_set_transport_tracers): needed once before construct_fields."""
        ia = lambda v: np.ascontiguousarray(v, dtype=np.int32)
        tt, dp, hd = ia(tracer_type), ia(depend), ia(has_dependents)
        if not (tt.size == dp.size == hd.size):
            raise ValueError("tracer_type, depend, has_dependents differ in length")
        geo = [np.ascontiguousarray(f[k], dtype=np.float64) for k in ("hm", "xav", "yav", "xxav", "yyav")]
        for g in geo:
            if g.shape != self.shape:
                raise ValueError(f"transport geometry plane has shape {g.shape}, block is {self.shape}")
        self._check(self.lib.evp_gpu_set_transport_tracers(self._h, C.c_int(int(ncat)), C.c_int(tt.size), _ip(tt), _ip(dp),
                                                           _ip(hd), *[_dp(g) for g in geo]))
        self._rc_shape = (int(ncat), int(tt.size))

    def transport_timing(self):
        """Device ms of the last departure_points (kernels, ghost updates) and construct_fields (kernel, ghost updates)."""
        ms = (C.c_double * 6)()
        self._check(self.lib.evp_gpu_get_transport_timing(self._h, ms))
        return {"departure_ms": ms[0], "departure_halo_ms": ms[1], "reconstruct_ms": ms[2], "reconstruct_halo_ms": ms[3],
                "flux_update_ms": ms[4], "remap_ms": ms[5]}

    def horizontal_remap_compute(self, dt):
        self._check(self.lib.evp_gpu_horizontal_remap_compute(self._h, C.c_double(dt)))

    def horizontal_remap_fetch(self):
        """Returns (mm, tm, (stop, category, istop, jstop)) of the last horizontal_remap."""
        ncat, nt = self._rc_shape
        mm, tm = np.zeros((ncat + 1,) + self.shape), np.zeros((ncat, nt) + self.shape)
        st, sc, i_s, j_s = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        self._check(self.lib.evp_gpu_horizontal_remap_fetch(self._h, _dp(mm), _dp(tm) if nt else None, C.byref(st), C.byref(sc),
                                                            C.byref(i_s), C.byref(j_s)))
        return mm, tm, (st.value, sc.value, i_s.value, j_s.value)

    def horizontal_remap(self, dt, mm, tm):
        """horizontal_remap (source/ice_transport_remap.F90:309-850) of this block from the resident velocities: mm
        (ncat+1, ny_block, nx_block), tm (ncat, ntrace, ny_block, nx_block) with ghost cells.  Returns the new (mm, tm,
        (stop, category, istop, jstop)); ghost cells of the results are not updated (the caller's halo update follows)."""
        self.construct_fields_upload(mm, tm)
        self.horizontal_remap_compute(dt)
        return self.horizontal_remap_fetch()

    # ---- transport_remap: from and to the ice state (aice0, aicen, trcrn, vicen, vsnon) ----
    def transport_remap_upload(self, st):
        ncat, nt = self._rc_shape
        a = {k: np.ascontiguousarray(st[k], dtype=np.float64) for k in ("aice0", "aicen", "trcrn", "vicen", "vsnon")}
        if a["aice0"].shape != self.shape or a["aicen"].shape != (ncat,) + self.shape or \
                a["trcrn"].shape != (ncat, nt - 2) + self.shape:
            raise ValueError("state arrays do not match ncat, ntrace - 2 and the block")
        self._check(self.lib.evp_gpu_transport_remap_upload(self._h, _dp(a["aice0"]), _dp(a["aicen"]),
                                                            _dp(a["trcrn"]) if nt > 2 else None, _dp(a["vicen"]), _dp(a["vsnon"])))

    def transport_remap_compute(self, dt, nt_qsno, nslyr, rhos, Lfresh):
        self._check(self.lib.evp_gpu_transport_remap_compute(self._h, C.c_double(dt), C.c_int(int(nt_qsno)), C.c_int(int(nslyr)),
                                                             C.c_double(rhos), C.c_double(Lfresh)))

    def transport_remap_fetch(self):
        ncat, nt = self._rc_shape
        o = {"aice0": np.zeros(self.shape), "aicen": np.zeros((ncat,) + self.shape), "trcrn": np.zeros((ncat, nt - 2) + self.shape),
             "vicen": np.zeros((ncat,) + self.shape), "vsnon": np.zeros((ncat,) + self.shape)}
        st, sc, i_s, j_s = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        self._check(self.lib.evp_gpu_transport_remap_fetch(self._h, _dp(o["aice0"]), _dp(o["aicen"]),
                                                           _dp(o["trcrn"]) if nt > 2 else None, _dp(o["vicen"]), _dp(o["vsnon"]),
                                                           C.byref(st), C.byref(sc), C.byref(i_s), C.byref(j_s)))
        return o, (st.value, sc.value, i_s.value, j_s.value)

    def transport_remap(self, dt, st, nt_qsno, nslyr, rhos, Lfresh):
        """transport_remap (source/ice_transport_driver.F90:198-627) of this block from the resident velocities.  st: dict
        aice0, aicen, trcrn, vicen, vsnon.  Returns (new state dict, (stop, category, istop, jstop))."""
        self.transport_remap_upload(st)
        self.transport_remap_compute(dt, nt_qsno, nslyr, rhos, Lfresh)
        return self.transport_remap_fetch()

    def construct_fields_compute(self):
        self._check(self.lib.evp_gpu_construct_fields_compute(self._h))

    def construct_fields_upload(self, mm, tm):
        ncat, nt = self._rc_shape
        mm, tm = np.ascontiguousarray(mm, dtype=np.float64), np.ascontiguousarray(tm, dtype=np.float64)
        if mm.shape != (ncat + 1,) + self.shape or tm.shape != (ncat, nt) + self.shape:
            raise ValueError(f"mm {mm.shape} / tm {tm.shape} do not match ncat = {ncat}, ntrace = {nt}, block {self.shape}")
        self._check(self.lib.evp_gpu_construct_fields_upload(self._h, _dp(mm), _dp(tm) if nt else None))

    def construct_fields_fetch(self):
        ncat, nt = self._rc_shape
        o = {k: np.zeros((ncat + 1,) + self.shape) for k in ("mc", "mx", "my")}
        o.update({k: np.zeros((ncat, nt) + self.shape) for k in ("tc", "tx", "ty")})
        self._check(self.lib.evp_gpu_construct_fields_fetch(self._h, *[_dp(o[k]) if o[k].size else None
                                                                       for k in ("mc", "mx", "my", "tc", "tx", "ty")]))
        return o

    def construct_fields(self, mm, tm):
        """make_masks + construct_fields + limited_gradient (source/ice_transport_remap.F90:867-1484) for categories
        0..ncat and the ghost updates that follow (:579-584, :616-621).  mm: (ncat+1, ny_block, nx_block); tm: (ncat,
        ntrace, ny_block, nx_block).  Returns a dict mc, mx, my, tc, tx, ty."""
        self.construct_fields_upload(mm, tm)
        self._check(self.lib.evp_gpu_construct_fields_compute(self._h))
        return self.construct_fields_fetch()

    # -- EAP (kdyn = 2)
    def set_eap_tables(self, tables: dict):
        """init_eap's lookup tables s11r ... s22s, each (21, 41, 41) = Fortran (nx_yield, ny_yield, na_yield)."""
        for n in EAP_TABLES:
            if tables[n].shape != EAP_TABLE_SHAPE:
                raise EvpError(f"table {n} has shape {tables[n].shape}, expected {EAP_TABLE_SHAPE}")
        t = EapTablesS(**{n: _dp(tables[n]) for n in EAP_TABLES})
        self._check(self.lib.evp_gpu_set_eap_tables(self._h, C.byref(t)))

    @staticmethod
    def _eap_state_struct(eap):
        st = EapStateS()
        for c in range(4):
            if eap.get(f"a11_{c + 1}") is not None:
                st.a11[c] = _dp(eap[f"a11_{c + 1}"])
            if eap.get(f"a12_{c + 1}") is not None:
                st.a12[c] = _dp(eap[f"a12_{c + 1}"])
        if eap.get("a11") is not None:
            st.a11_avg = _dp(eap["a11"])
        if eap.get("a12") is not None:
            st.a12_avg = _dp(eap["a12"])
        return st

    def upload_eap(self, eap: dict):
        """The structure tensor a11_1..4, a12_1..4 (and the averages a11, a12) of this block; missing keys stay."""
        st = self._eap_state_struct(eap)
        self._check(self.lib.evp_gpu_upload_eap(self._h, C.byref(st)))

    def download_eap(self, eap: dict):
        """Fills the arrays of `eap` that are present: structure tensor, averages and the EAP_OUT diagnostics."""
        st = self._eap_state_struct(eap)
        out = EapOutS(**{n: _dp(eap[n]) for n in EAP_OUT if eap.get(n) is not None})
        self._check(self.lib.evp_gpu_download_eap(self._h, C.byref(st), C.byref(out)))

    def stats(self) -> Stats:
        st = Stats()
        self._check(self.lib.evp_gpu_get_stats(self._h, C.byref(st)))
        return st

    def principal_stress(self):
        """sig1, sig2 (principal_stress, source/ice_dyn_shared.F90:853-893) of the resident state."""
        s1, s2 = np.empty(self.shape), np.empty(self.shape)
        self._check(self.lib.evp_gpu_principal_stress(self._h, _dp(s1), _dp(s2)))
        return s1, s2

    def trace(self):
        """(ndte, 8) uint64 device timestamps of the last compute (needs EVP_B200_TRACE=1)."""
        buf = np.zeros((self.cfg.ndte, 8), dtype=np.uint64)
        self._check(self.lib.evp_gpu_get_trace(self._h, buf.ctypes.data_as(C.POINTER(C.c_uint64)),
                                               C.c_int(buf.size)))
        return buf

    def synchronize(self):
        self._check(self.lib.evp_gpu_synchronize(self._h))

    def peer_blob(self) -> bytes:
        n = self.lib.evp_gpu_peer_blob_size()
        buf = C.create_string_buffer(n)
        self._check(self.lib.evp_gpu_get_peer_blob(self._h, buf))
        return buf.raw

    def set_peers(self, blobs: list):
        raw = b"".join(blobs)
        self._check(self.lib.evp_gpu_set_peers(self._h, raw))

    def close(self):
        if self._h:
            self.lib.evp_gpu_finalize(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def compute_group(solvers: list, dt: float):
    """evp_gpu_compute_group: all ranks of one decomposition driven from this thread."""
    arr = (C.c_void_p * len(solvers))(*[e._h for e in solvers])
    solvers[0]._check(solvers[0].lib.evp_gpu_compute_group(arr, len(solvers), C.c_double(dt)))


def principal_stress_static(stressp_1, stressm_1, stress12_1, prs_sig, device=0, strict=True):
    """evp_gpu_principal_stress_static: sig1, sig2 from host arrays, no handle, no allocation on the device (the scratch
    is a static device array of STATIC_MAX_CELLS cells per plane).  Raises EvpError beyond that size."""
    lib = load_library(strict)
    ny, nx = stressp_1.shape
    sig1, sig2 = np.empty_like(stressp_1), np.empty_like(stressp_1)
    rc = lib.evp_gpu_principal_stress_static(C.c_int(device), C.c_int(nx), C.c_int(ny), _dp(stressp_1), _dp(stressm_1),
                                             _dp(stress12_1), _dp(prs_sig), _dp(sig1), _dp(sig2))
    if rc != 0:
        msg = lib.evp_gpu_strerror(rc).decode()
        if rc in (3, 7):
            msg += " [" + lib.evp_gpu_last_cuda_error().decode() + "]"
        raise EvpError(msg)
    return sig1, sig2


def device_count() -> int:
    return load_library().evp_gpu_device_count()


def run_case(case, strict=True, **kw):
    """One evp(dt) call on a fixtures.Case through the C ABI (single GPU). Returns Stats."""
    e = EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=strict, **kw)
    try:
        e.set_grid(case.grid.f)
        e.evp(case.dt, case.inp, case.io, case.out)
        return e.stats()
    finally:
        e.close()


def run_case_eap(case, eap, tables, strict=True, **kw):
    """One eap(dt) call (kdyn = 2) on a fixtures.Case and an EAP state dict (structure tensor in, every EAP field out)."""
    e = EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=strict, kdyn=2, **kw)
    try:
        e.set_grid(case.grid.f)
        e.set_eap_tables(tables)
        e.upload_eap(eap)
        e.evp(case.dt, case.inp, case.io, case.out)
        e.download_eap(eap)
        return e.stats()
    finally:
        e.close()


class EvpGroup:
    """All x-slabs of one decomposition driven from one process (one device): the host-side mirror
    of `nranks` tasks each calling evp(dt) on its own block.  Global-with-ghost arrays in and out."""

    def __init__(self, dom, ndte, dt_init, nranks, device=0, slab_start=None, **kw):
        from . import fixtures as fx
        self._fx = fx
        self.n = nranks
        self.starts = slab_start            # None = equal slabs
        self.ranks = [EvpGpu.for_domain(dom, ndte, dt_init, rank=r, nranks=nranks, device=device,
                                        slab_start=slab_start, **kw)
                      for r in range(nranks)]
        blobs = [e.peer_blob() for e in self.ranks]
        for e in self.ranks:
            e.set_peers(blobs)

    def _slabs(self, d, r):
        return {k: self._fx.slab(v, r, self.n, self.starts) for k, v in d.items() if v is not None}

    def set_grid(self, f):
        names = GRID_R8 + ("tmask", "umask") + tuple(k for k in ("HTE", "HTN") if f.get(k) is not None)
        for r, e in enumerate(self.ranks):
            e.set_grid(self._slabs({k: f[k] for k in names}, r))

    def evp(self, dt, inp, io, out):
        parts = []
        for r, e in enumerate(self.ranks):
            si, sio, so = self._slabs(inp, r), self._slabs(io, r), self._slabs(out, r)
            parts.append((sio, so))
            e.upload(si, sio)
        compute_group(self.ranks, dt)
        for e, (sio, so) in zip(self.ranks, parts):
            e.download(sio, so)
        for k in io:
            io[k][...] = self._fx.unslab([p[0][k] for p in parts])
        for k in out:
            if out[k] is not None:
                out[k][...] = self._fx.unslab([p[1][k] for p in parts])
        return [e.stats() for e in self.ranks]

    def upload(self, inp, io):
        for r, e in enumerate(self.ranks):
            e.upload(self._slabs(inp, r), self._slabs(io, r))

    def download(self, io, out):
        parts = []
        for r, e in enumerate(self.ranks):
            sio, so = self._slabs(io, r), self._slabs(out, r)
            e.download(sio, so)
            parts.append((sio, so))
        for k in io:
            io[k][...] = self._fx.unslab([p[0][k] for p in parts])
        for k in out:
            if out[k] is not None:
                out[k][...] = self._fx.unslab([p[1][k] for p in parts])

    def set_eap_tables(self, tables):
        for e in self.ranks:
            e.set_eap_tables(tables)

    def upload_eap(self, eap):
        for r, e in enumerate(self.ranks):
            e.upload_eap(self._slabs(eap, r))

    def download_eap(self, eap):
        parts = []
        for r, e in enumerate(self.ranks):
            se = self._slabs(eap, r)
            e.download_eap(se)
            parts.append(se)
        for k in eap:
            eap[k][...] = self._fx.unslab([p[k] for p in parts])

    def set_transport_grid(self, f):
        for r, e in enumerate(self.ranks):
            e.set_transport_grid(self._slabs({k: f[k] for k in ("dxu", "dyu", "HTN", "HTE")}, r))

    def departure_points(self, dt, l_dp_midpt=True):
        """All ranks; returns the global dpx, dpy and the per-rank (l_stop, istop, jstop)."""
        arr = (C.c_void_p * self.n)(*[e._h for e in self.ranks])
        self.ranks[0]._check(self.ranks[0].lib.evp_gpu_departure_points_group(arr, self.n, C.c_double(dt),
                                                                              C.c_int(int(l_dp_midpt))))
        parts = [e.departure_points_fetch() for e in self.ranks]
        return (self._fx.unslab([p[0] for p in parts]), self._fx.unslab([p[1] for p in parts]),
                [p[2:] for p in parts])

    def restart_fixup(self):
        arr = (C.c_void_p * self.n)(*[e._h for e in self.ranks])
        self.ranks[0]._check(self.ranks[0].lib.evp_gpu_restart_fixup_group(arr, self.n))

    def set_transport_tracers(self, ncat, tracer_type, depend, has_dependents, f):
        for r, e in enumerate(self.ranks):
            e.set_transport_tracers(ncat, tracer_type, depend, has_dependents,
                                    self._slabs({k: f[k] for k in ("hm", "xav", "yav", "xxav", "yyav")}, r))

    def transport_remap(self, dt, st, nt_qsno, nslyr, rhos, Lfresh):
        """All ranks; returns the global new state dict and the per-rank (stop, category, istop, jstop)."""
        for r, e in enumerate(self.ranks):
            e.transport_remap_upload({k: self._fx.slab(np.ascontiguousarray(v), r, self.n, self.starts) for k, v in st.items()})
        arr = (C.c_void_p * self.n)(*[e._h for e in self.ranks])
        self.ranks[0]._check(self.ranks[0].lib.evp_gpu_transport_remap_group(arr, self.n, C.c_double(dt), C.c_int(int(nt_qsno)),
                                                                             C.c_int(int(nslyr)), C.c_double(rhos),
                                                                             C.c_double(Lfresh)))
        parts = [e.transport_remap_fetch() for e in self.ranks]
        return {k: self._fx.unslab([p[0][k] for p in parts]) for k in parts[0][0]}, [p[1] for p in parts]

    def horizontal_remap(self, dt, mm, tm):
        """All ranks; returns the global new mm, tm and the per-rank (stop, category, istop, jstop)."""
        for r, e in enumerate(self.ranks):
            e.construct_fields_upload(self._fx.slab(mm, r, self.n, self.starts), self._fx.slab(tm, r, self.n, self.starts))
        arr = (C.c_void_p * self.n)(*[e._h for e in self.ranks])
        self.ranks[0]._check(self.ranks[0].lib.evp_gpu_horizontal_remap_group(arr, self.n, C.c_double(dt)))
        parts = [e.horizontal_remap_fetch() for e in self.ranks]
        return (self._fx.unslab([p[0] for p in parts]), self._fx.unslab([p[1] for p in parts]), [p[2] for p in parts])

    def construct_fields(self, mm, tm):
        """All ranks; returns the global mc, mx, my, tc, tx, ty."""
        for r, e in enumerate(self.ranks):
            e.construct_fields_upload(self._fx.slab(mm, r, self.n, self.starts), self._fx.slab(tm, r, self.n, self.starts))
        arr = (C.c_void_p * self.n)(*[e._h for e in self.ranks])
        self.ranks[0]._check(self.ranks[0].lib.evp_gpu_construct_fields_group(arr, self.n))
        parts = [e.construct_fields_fetch() for e in self.ranks]
        return {k: self._fx.unslab([p[k] for p in parts]) for k in parts[0]}

    def close(self):
        for e in self.ranks:
            e.close()
