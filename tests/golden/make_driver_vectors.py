#!/usr/bin/env python
"""Golden vectors for the WHOLE evp(dt) call from the reference's own text (build container only).

The driver itself -- subroutine evp, source/ice_dyn_evp.F90:68-510 -- is executed mechanically by
tests/golden/fortran_exec.py on a domain cut into several blocks, calling the reference's own

  source/ice_dyn_shared.F90  set_evp_parameters, evp_prep1, evp_prep2, stepu, evp_finish
  source/ice_dyn_evp.F90     evp, stress
  source/ice_grid.F90        to_ugrid, t2ugrid_vector, to_tgrid, u2tgrid_vector
  source/ice_mechred.F90     ice_strength, asum_ridging, ridge_itd
  source/ice_blocks.F90      create_blocks, ice_blocksGetNbrID, get_block_parameter, get_block
  mpi/ice_boundary.F90       ice_HaloCreate, ice_HaloMsgCreate, ice_HaloUpdate2DR8 / 3DR8 / 2DI4 / _stress

Stood in for: the generic name ice_HaloUpdate (dispatch on rank and type to the specific routines above,
as the Fortran interface block does), ice_distributionGet* (one task owns every block), global_minval /
global_maxval (one task), timers (no-ops).  Inputs come from cice5_b200.fixtures (grid with its ghost
ring, ice state, forcing, 5-category thickness distribution), scattered to the blocks as windows of the
global arrays-with-ghost-ring.  Every inout / out field is gathered back to a global array-with-ghost-ring
and stored in tests/golden/ref_driver_vectors.npz together with the inputs;
tests/test_oracle_reference_text.py runs the C oracle (halos on, 1 / 2 / 4 slabs) on the same inputs
and demands bitwise equality on the physical cells and on the ring cells the reference defines.

    python tests/golden/make_driver_vectors.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import fortran_exec as fe  # noqa: E402
from cice5_b200 import fixtures as fx  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(HERE, "ref_driver_vectors.npz")
NCAT = 5

# name: (grid factory, block_size_x, block_size_y, CPP macros, revised_evp, ndte, scenario)
CASES = {
    "tripole_2x2": (lambda: fx.synthetic_tripole(24, 12), 12, 6, (), False, 4, "full"),
    "tripole_wet_4x1": (lambda: fx.synthetic_tripole(24, 12, wet_poles=True), 6, 12, (), False, 3, "full"),
    "tripole_coupled_1x1": (lambda: fx.synthetic_tripole(24, 12), 24, 12, ("coupled",), False, 3, "caps"),
    "open_3x2": (lambda: fx.synthetic_displaced(18, 12), 6, 6, (), False, 4, "full"),
    "open_revised_2x1": (lambda: fx.synthetic_displaced(18, 12), 9, 12, (), True, 3, "full"),
    # a regional basin: E-W and N-S closed, ocean up to three of its edges
    "closed_2x2": (lambda: fx.synthetic_closed(16, 12), 8, 6, (), False, 4, "full"),
    # the ACCESS build with calc_strair = .false. (ice_dyn_evp.F90:220-245): strairx/y are the coupler's
    # U-grid stresses strax/stray, t2ugrid_vector is skipped
    "tripole_access_2x1": (lambda: fx.synthetic_tripole(24, 12), 12, 12, ("ACCESS",), False, 3, "full"),
    # the ACCESS-OM production build, bld/Macros.nci:72,81: -DAusCOM -Dcoupled -DACCESS, with the namelist's ocean
    # turning angle (cosw, sinw are variables there, ice_dyn_shared.F90:66-72) and use_ocnslope = .false.
    # (drivers/auscom/cpl_parameters.F90:31: the tilt comes from the currents although `coupled` is defined) ...
    "tripole_auscom_2x2": (lambda: fx.synthetic_tripole(24, 12), 12, 6, ("AusCOM", "coupled", "ACCESS"), False, 3, "full",
                           {"use_ocnslope": False, "angle_deg": 25.0}),
    # ... and with use_ocnslope = .true. (tilt from the sea-surface slope)
    "tripole_auscom_slope_1x2": (lambda: fx.synthetic_tripole(24, 12), 24, 6, ("AusCOM", "coupled", "ACCESS"), False, 3,
                                 "caps", {"use_ocnslope": True, "angle_deg": 0.0}),
}
# The reference's real gx3 grid and the 1-degree tripole: too large to commit field by field -- their inputs
# are regenerated from cice5_b200.fixtures (deterministic) and only SHA-256 digests of the inputs and of every
# output block array go to tests/golden/ref_driver_digests.json (`--big`, a few minutes of interpreter time).
BIG_CASES = {
    "gx3_real_1x1": (lambda: fx.make_grid("gx3"), 100, 116, (), False, 3, "full"),
    "gx3_real_2x2": (lambda: fx.make_grid("gx3"), 50, 58, (), False, 3, "full"),
    "tp1_4x2": (lambda: fx.make_grid("tp1"), 90, 150, (), False, 2, "caps"),
    # BASELINE.json's "ACCESS-OM 1 degree tripole 360x300" in its production build (bld/Macros.nci:72,81)
    "tp1_auscom_4x2": (lambda: fx.make_grid("tp1"), 90, 150, ("AusCOM", "coupled", "ACCESS"), False, 2, "caps",
                       {"use_ocnslope": False, "angle_deg": 25.0}),
}
DIGESTS = os.path.join(HERE, "ref_driver_digests.json")

STATE_R8 = ("aice", "vice", "vsno", "aice_init", "aice0", "strairxt", "strairyt", "uocn", "vocn", "ss_tltx",
            "ss_tlty", "cdn_ocn", "strength", "uvel", "vvel", "fm", "strairx", "strairy", "strtltx", "strtlty",
            "strocnx", "strocny", "strintx", "strinty", "strocnxt", "strocnyt", "divu", "shear", "rdg_conv",
            "rdg_shear", "prs_sig", "uvel_init", "vvel_init", "strax", "stray") + \
    tuple(f"stress{k}_{c}" for k in ("p", "m", "12") for c in (1, 2, 3, 4))
GRID_R8 = ("dxt", "dyt", "dxhy", "dyhx", "cxp", "cyp", "cxm", "cym", "tarea", "uarea", "tarear", "uarear", "tinyarea")


def _lines(rel, defined=()):
    return fe.preprocess(open(os.path.join(REF, rel)).read(), defined)


def machine(dom, bsx, bsy, macros, revised_evp, ndte, extra=None):
    ew = "cyclic" if dom.ew_cyclic else "closed"
    ns = "tripole" if dom.ns_tripole else ("open" if dom.ns_open else "closed")
    nsd = {"block_size_x": bsx, "block_size_y": bsy}
    fe.parse_parameters(_lines("drivers/cice/ice_constants.F90"), nsd)
    src = {k: _lines(v, macros) for k, v in (("blocks", "source/ice_blocks.F90"), ("bnd", "mpi/ice_boundary.F90"),
                                             ("shared", "source/ice_dyn_shared.F90"), ("evp", "source/ice_dyn_evp.F90"),
                                             ("grid", "source/ice_grid.F90"), ("mech", "source/ice_mechred.F90"))}
    for k in ("blocks", "bnd", "shared", "mech"):
        fe.parse_parameters(src[k], nsd)
    m = fe.Module(nsd)
    g = m.ns
    if "AusCOM" in macros:          # namelist variables of that build (ice_init.F90:167; cpl_parameters.F90:31)
        import math
        a = math.radians(extra["angle_deg"])
        g.update(cosw=math.cos(a), sinw=math.sin(a), use_ocnslope=extra["use_ocnslope"])

    @fe.fstub(["dist", "nprocs", "communicator", "numlocalblocks"], out=["nprocs", "communicator", "numlocalblocks"])
    def ice_distributionget(dist, nprocs=None, communicator=None, numlocalblocks=None):
        return {"nprocs": 1, "communicator": 0, "numlocalblocks": g["nblocks_tot"]}

    @fe.fstub(["dist", "blockid", "processor", "localid"], out=["processor", "localid"])
    def ice_distributiongetblockloc(dist, blockid, processor, localid):
        return {"processor": 1, "localid": blockid}

    def ice_haloupdate(array, halo, fieldloc, fieldkind, fillvalue=None):
        """The generic interface (mpi/ice_boundary.F90:65-76): specific routine by rank and type."""
        kw = {} if fillvalue is None else {"fillvalue": fillvalue}
        integer = array.a.dtype.kind in "iu" or (array.a.dtype == object and isinstance(array.a.flat[0], int))
        name = {(3, False): "ice_haloupdate2dr8", (3, True): "ice_haloupdate2di4", (4, False): "ice_haloupdate3dr8"}[
            (array.a.ndim, bool(integer))]
        return g[name](array, halo, fieldloc, fieldkind, **kw)

    def global_minval(arr, dist, mask):
        vals = [arr(i, j, n) for n in range(1, g["nblocks"] + 1)
                for j in range(g["all_blocks"](n).jlo, g["all_blocks"](n).jhi + 1)
                for i in range(g["all_blocks"](n).ilo, g["all_blocks"](n).ihi + 1) if mask(i, j, n)]
        return min(vals)

    g.update(ice_distributionget=ice_distributionget, ice_distributiongetblockloc=ice_distributiongetblockloc,
             ice_haloupdate=ice_haloupdate, global_maxval=lambda x, dist: x, global_minval=global_minval,
             my_task=0, master_task=0, nu_diag=6, mpi_status_size=6, distrb_info=None,
             all_blocks=None, all_blocks_ij=None, i_global=None, j_global=None, nblocks_x=0, nblocks_y=0,
             nblocks_tot=0, buftripoler8=None, buftripolei4=None, buftripoler4=None, bufsizesend=0, bufsizerecv=0,
             bufsendr8=None, bufrecvr8=None, bufsendi4=None, bufrecvi4=None, bufsendr4=None, bufrecvr4=None,
             ndte=ndte, revised_evp=revised_evp, maskhalo_dyn=False, ncat=NCAT, kstrength=1, krdg_partic=1,
             krdg_redist=1, mu_rdg=3.0, cf=17.0, calc_strair="ACCESS" not in macros, grid_type="tripole" if dom.grid_tripole else "displaced_pole")
    m.load(src["blocks"], ["get_block_parameter", "create_blocks", "ice_blocksgetnbrid", "get_block"],
           derived_arrays=["all_blocks"])
    m.load(src["bnd"], ["ice_haloincrementmsgcount", "ice_halomsgcreate", "ice_halocreate", "ice_haloupdate2dr8",
                        "ice_haloupdate3dr8", "ice_haloupdate2di4", "ice_haloupdate_stress"])
    m.call("create_blocks", nx_global=dom.nx, ny_global=dom.ny, ew_boundary_type=ew, ns_boundary_type=ns)
    g["halo_info"] = m.call("ice_halocreate", dist=None, nsboundarytype=ns, ewboundarytype=ew, nxglobal=dom.nx)
    nb = g["nblocks_tot"]
    g.update(nblocks=nb, max_blocks=nb, blocks_ice=fe.FArr.from_numpy(np.arange(1, nb + 1)))
    m.load(src["grid"], ["to_ugrid", "t2ugrid_vector", "to_tgrid", "u2tgrid_vector"], callable_names=["ice_haloupdate"])
    m.load(src["mech"], ["asum_ridging", "ridge_itd", "ice_strength"])
    m.load(src["shared"], ["set_evp_parameters", "evp_prep1", "evp_prep2", "stepu", "evp_finish"])
    m.load(src["evp"], ["stress", "evp"], callable_names=["ice_haloupdate"])
    return m


def windows(m, dom, bsx, bsy):
    """[(n, row slice, column slice)]: block n (1-based) as a window of a global array-with-ghost-ring."""
    out = []
    for n in range(1, m.ns["nblocks_tot"] + 1):
        b = m.ns["all_blocks"](n)
        r0, c0 = (b.jblock - 1) * bsy, (b.iblock - 1) * bsx
        out.append((n, slice(r0, r0 + bsy + 2), slice(c0, c0 + bsx + 2)))
    return out


def scatter(m, wins, glob, dtype=float):
    """Global (ny+2, nx+2) -> Fortran (nx_block, ny_block, nblocks), every block = its window, ghosts included."""
    arr = fe.FArr([(1, m.ns["nx_block"]), (1, m.ns["ny_block"]), (1, len(wins))], dtype=dtype)
    for n, rs, cs in wins:
        arr.a[:, :, n - 1] = np.transpose(glob[rs, cs])
    return arr


def gather(m, wins, arr, shape, dtype=float):
    """Back to a global (ny+2, nx+2) array: physical cells from their owner, ring cells from the block whose
    window holds them (a ring cell seen by two blocks is taken from the first; the test compares per block)."""
    out = np.zeros(shape, dtype=dtype)
    seen = np.zeros(shape, bool)
    for n, rs, cs in wins:                                          # ring / ghost cells first
        blk = np.transpose(arr.a[:, :, n - 1])
        sub, ss = out[rs, cs], seen[rs, cs]
        sub[~ss] = blk[~ss]
        ss[...] = True
    for n, rs, cs in wins:                                          # then every block's physical cells
        blk = np.transpose(arr.a[:, :, n - 1])
        out[rs, cs][1:-1, 1:-1] = blk[1:-1, 1:-1]
    return out


def blocks_of(arr):
    return np.ascontiguousarray(np.transpose(np.asarray(arr.a, dtype=float), (2, 1, 0)))


def _digest(a) -> str:
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main(out_path=OUT, big=False):
    pack = {}
    digests = {}
    for cname, spec in (BIG_CASES if big else CASES).items():
        factory, bsx, bsy, macros, revised, ndte, scenario = spec[:7]
        extra = spec[7] if len(spec) > 7 else None
        grid = factory()
        dom = grid.dom
        case = fx.make_case(grid, scenario, ndte=ndte, dt=3600.0)
        itd = fx.ice_itd(case.inp)
        m = machine(dom, bsx, bsy, macros, revised, ndte, extra)
        g = m.ns
        wins = windows(m, dom, bsx, bsy)
        shape = (dom.ny + 2, dom.nx + 2)
        zero = np.zeros(shape)
        # --- module arrays of ice_grid / ice_state / ice_flux, block layout
        for n in GRID_R8:
            g[n] = scatter(m, wins, grid.f[n])
        g["fcor_blk"] = scatter(m, wins, grid.f["fcor"])
        g["tmask"] = scatter(m, wins, grid.f["tmask"] > 0, dtype=bool)
        g["umask"] = scatter(m, wins, grid.f["umask"] > 0, dtype=bool)
        inp, io = case.inp, case.io
        src = {"aice": inp["aice"], "vice": inp["vice"], "vsno": inp["vsno"], "aice_init": inp["aice_init"],
               "aice0": itd["aice0"], "strairxt": inp["strairxT"], "strairyt": inp["strairyT"], "uocn": inp["uocn"],
               "vocn": inp["vocn"], "ss_tltx": inp["ss_tltx"], "ss_tlty": inp["ss_tlty"], "cdn_ocn": inp["Cw"],
               "uvel": io["uvel"], "vvel": io["vvel"], "fm": io["fm"]}
        src.update({k: io[k] for k in io if k.startswith("stress")})
        if "ACCESS" in macros:                     # wind stress from the coupler, already on the U grid
            inp = dict(inp, strax=0.7 * inp["strairyT"] + 0.01, stray=-1.3 * inp["strairxT"])
            src.update(strax=inp["strax"], stray=inp["stray"])
        for n in STATE_R8 + (("sicemass",) if "AusCOM" in macros else ()):
            g[n] = scatter(m, wins, src.get(n, zero))
        g["iceumask"] = scatter(m, wins, io["iceumask"] > 0, dtype=bool)
        for n4, a4 in (("aicen", itd["aicen"]), ("vicen", itd["vicen"])):
            t = fe.FArr([(1, g["nx_block"]), (1, g["ny_block"]), (1, NCAT), (1, len(wins))])
            for k in range(NCAT):
                t.a[:, :, k, :] = scatter(m, wins, a4[k]).a
            g[n4] = t
        # --- init_evp's share (set_evp_parameters), then the driver
        m.call("set_evp_parameters", dt=case.dt)
        m.call("evp", dt=case.dt)
        outs = {"uvel": "uvel", "vvel": "vvel", "fm": "fm", "strairx": "strairx", "strairy": "strairy",
                "strtltx": "strtltx", "strtlty": "strtlty", "strocnx": "strocnx", "strocny": "strocny",
                "strintx": "strintx", "strinty": "strinty", "strocnxT": "strocnxt", "strocnyT": "strocnyt",
                "divu": "divu", "shear": "shear", "rdg_conv": "rdg_conv", "rdg_shear": "rdg_shear",
                "prs_sig": "prs_sig", "uvel_init": "uvel_init", "vvel_init": "vvel_init", "strength": "strength"}
        outs.update({k: k for k in io if k.startswith("stress")})
        if big:                                    # digests only; physical cells and whole blocks separately
            d = {"dims": [dom.nx, dom.ny, bsx, bsy, int(g["nblocks_x"]), int(g["nblocks_y"]), ndte],
                 "scenario": scenario, "dt": case.dt, "access": "ACCESS" in macros,
                 "kw": {"cosw": float(g["cosw"]), "sinw": float(g["sinw"]),
                        "tilt_mode": int("coupled" in macros and not ("AusCOM" in macros and not extra["use_ocnslope"]))},
                 "inputs": {n: _digest(v) for n, v in list(inp.items()) + list(io.items()) + list(itd.items())},
                 "grid": {n: _digest(grid.f[n]) for n in fx.GRID_FIELDS + ("tmask", "umask")},
                 "physical": {}, "blocks": {}}
            for n, fn in list(outs.items()) + [("iceumask", "iceumask")]:
                blk = blocks_of(g[fn])
                d["physical"][n] = _digest(blk[:, 1:-1, 1:-1])
                d["blocks"][n] = _digest(blk)
            digests[cname] = d
            print(cname, "done", flush=True)
            continue
        # --- store
        key = cname + "__"
        pack[key + "dims"] = np.array([dom.nx, dom.ny, bsx, bsy, g["nblocks_x"], g["nblocks_y"], ndte], np.int64)
        pack[key + "flags"] = np.array([int(dom.ew_cyclic), int(dom.ns_tripole), int(dom.grid_tripole),
                                        int("coupled" in macros and not ("AusCOM" in macros and not extra["use_ocnslope"])),
                                        int(revised), int("ACCESS" in macros),
                                        int(not dom.ns_tripole and not dom.ns_open)], np.int64)
        pack[key + "angle"] = np.array([g["cosw"], g["sinw"]], float)
        pack[key + "dt"] = np.float64(case.dt)
        for n in fx.GRID_FIELDS + ("tmask", "umask"):
            pack[key + "grid__" + n] = grid.f[n]
        for n, v in inp.items():
            pack[key + "in__" + n] = v
        for n, v in io.items():
            pack[key + "io__" + n] = v
        for n, v in itd.items():
            pack[key + "itd__" + n] = v
        for n, fn in outs.items():
            pack[key + "out__" + n] = blocks_of(g[fn])
        pack[key + "out__iceumask"] = blocks_of(g["iceumask"]).astype(np.int32)
        pack[key + "consts"] = np.array([g["arlx1i"], g["brlx"], g["denom1"], g["revp"], g["ecci"]], float)
    if fe.UNTRANSLATED:
        print("statements outside the interpreter's subset (raise if reached):")
        for u_ in sorted(set(fe.UNTRANSLATED)):
            print("   ", u_)
    if big:
        import json
        with open(DIGESTS, "w") as fh:
            json.dump(digests, fh, indent=1, sort_keys=True)
        print(f"{DIGESTS}: {len(digests)} cases")
        return
    np.savez_compressed(out_path, **pack)
    print(f"{out_path}: {len(pack)} arrays, {os.path.getsize(out_path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main(big="--big" in sys.argv)
