"""Independent vectorised numpy restatement of CICE5 evp(dt) on ONE block (TEST INFRASTRUCTURE).

Written from the Fortran (not from evp_oracle.c) to cross-check the C oracle: the two must agree
BITWISE (numpy evaluates the same IEEE operations element-wise, no FMA contraction).  Only tests/
may import this.  Parity status: see oracle/evp_oracle.h.  Citations relative to the reference:
  source/ice_dyn_evp.F90:68-510 (evp), :520-849 (stress)
  source/ice_dyn_shared.F90:185-259, 270-365, 377-614, 623-748, 757-844
  source/ice_grid.F90:1799-1958 ; halos: cice5_b200.fixtures.halo_update (serial/ice_boundary.F90:630-896)
Arrays: (ny+2, nx+2) C-order == Fortran (nx_block, ny_block); a[j-1, i-1] is Fortran a(i,j).
"""
from __future__ import annotations

import numpy as np

from cice5_b200 import fixtures as fx


def _sh(a, di, dj):
    """View of `a` shifted so that result[j, i] = a[j+dj, i+di] on the physical window."""
    ny, nx = a.shape[0] - 2, a.shape[1] - 2
    return a[1 + dj:1 + dj + ny, 1 + di:1 + di + nx]


def evp_parameters(dt, ndte, revised_evp=False, xmin=None):
    """set_evp_parameters (ice_dyn_shared.F90:209-257); xmin = min(dxt, dyt) over tmask cells."""
    dte = dt / float(ndte)
    dtei = 1.0 / dte
    ecci = 0.25
    tdamp2 = 2.0 * 0.36 * dt
    dte2T = dte / tdamp2
    if revised_evp:                        # Bouillon et al. 2013 (:225-233)
        Se, xi = 0.86, 5.5e-3
        gamma = 0.25 * 1.0e11 * dt
        arlx1i = 2.0 * xi / Se
        brlx = 2.0 * Se * xi * gamma / (xmin * xmin)
    else:
        arlx1i = dte2T
        brlx = dt * dtei
    denom1 = 1.0 / (1.0 + arlx1i)
    return ecci, arlx1i, brlx, denom1


def to_ugrid(w, tarea, uarea):
    out = np.zeros_like(w)
    P = (slice(1, -1), slice(1, -1))
    out[P] = 0.25 * (_sh(w, 0, 0) * _sh(tarea, 0, 0) + _sh(w, 1, 0) * _sh(tarea, 1, 0)
                     + _sh(w, 0, 1) * _sh(tarea, 0, 1) + _sh(w, 1, 1) * _sh(tarea, 1, 1)) / _sh(uarea, 0, 0)
    return out


def to_tgrid(w, tarea, uarea, out):
    P = (slice(1, -1), slice(1, -1))
    out[P] = 0.25 * (_sh(w, 0, 0) * _sh(uarea, 0, 0) + _sh(w, -1, 0) * _sh(uarea, -1, 0)
                     + _sh(w, 0, -1) * _sh(uarea, 0, -1) + _sh(w, -1, -1) * _sh(uarea, -1, -1)) / _sh(tarea, 0, 0)
    return out


def stress(k, ksub, ndte, tact, u, v, g, strength, S, out):
    """ice_dyn_evp.F90:618-847 on T cells (2..nx+2, 2..ny+2) where `tact`; returns str[8]."""
    ecci, arlx1i, brlx, denom1 = k
    p111 = 1.0 / 9.0; p055 = p111 * 0.5; p027 = p055 * 0.5
    p166 = 1.0 / 6.0; p333 = 1.0 / 3.0; p222 = 2.0 / 9.0
    W = (slice(1, None), slice(1, None))            # T cells i = 2..nx+2, j = 2..ny+2 (0-based 1..)
    uij, vij = u[1:, 1:], v[1:, 1:]
    uwj, vwj = u[1:, :-1], v[1:, :-1]
    uis, vis = u[:-1, 1:], v[:-1, 1:]
    uws, vws = u[:-1, :-1], v[:-1, :-1]
    cxp, cyp, cxm, cym = (g[n][W] for n in ("cxp", "cyp", "cxm", "cym"))
    dxt, dyt, dxhy, dyhx = (g[n][W] for n in ("dxt", "dyt", "dxhy", "dyhx"))
    tarear, tiny = g["tarear"][W], g["tinyarea"][W]
    st = strength[W]
    divune = cyp * uij - dyt * uwj + cxp * vij - dxt * vis
    divunw = cym * uwj + dyt * uij + cxp * vwj - dxt * vws
    divusw = cym * uws + dyt * uis + cxm * vws + dxt * vwj
    divuse = cyp * uis - dyt * uws + cxm * vis + dxt * vij
    tenne = -cym * uij - dyt * uwj + cxm * vij + dxt * vis
    tennw = -cyp * uwj + dyt * uij + cxm * vwj + dxt * vws
    tensw = -cyp * uws + dyt * uis + cxp * vws - dxt * vwj
    tense = -cym * uis - dyt * uws + cxp * vis - dxt * vij
    shne = -cym * vij - dyt * vwj - cxm * uij - dxt * uis
    shnw = -cyp * vwj + dyt * vij - cxm * uwj - dxt * uws
    shsw = -cyp * vws + dyt * vis - cxp * uws + dxt * uwj
    shse = -cym * vis - dyt * vws - cxp * uis + dxt * uij
    Dne = np.sqrt(divune * divune + ecci * (tenne * tenne + shne * shne))
    Dnw = np.sqrt(divunw * divunw + ecci * (tennw * tennw + shnw * shnw))
    Dse = np.sqrt(divuse * divuse + ecci * (tense * tense + shse * shse))
    Dsw = np.sqrt(divusw * divusw + ecci * (tensw * tensw + shsw * shsw))
    m = tact[W]
    if ksub == ndte:
        divu = 0.25 * (divune + divunw + divuse + divusw) * tarear
        tmp = 0.25 * (Dne + Dnw + Dse + Dsw) * tarear
        ts = tenne + tennw + tense + tensw
        ss = shne + shnw + shse + shsw
        for name, val in (("divu", divu), ("rdg_conv", -np.minimum(divu, 0.0)),
                          ("rdg_shear", 0.5 * (tmp - np.abs(divu))),
                          ("shear", 0.25 * tarear * np.sqrt(ts * ts + ss * ss))):
            out[name][W] = np.where(m, val, out[name][W])
    with np.errstate(divide="ignore", invalid="ignore"):
        c0ne = st / np.maximum(Dne, tiny); c0nw = st / np.maximum(Dnw, tiny)
        c0sw = st / np.maximum(Dsw, tiny); c0se = st / np.maximum(Dse, tiny)
    out["prs_sig"][W] = np.where(m, c0ne * Dne, out["prs_sig"][W])
    c1ne, c1nw, c1sw, c1se = c0ne * arlx1i, c0nw * arlx1i, c0sw * arlx1i, c0se * arlx1i
    c0ne, c0nw, c0sw, c0se = c1ne * ecci, c1nw * ecci, c1sw * ecci, c1se * ecci
    new = {}
    new["stressp_1"] = (S["stressp_1"][W] + c1ne * (divune - Dne)) * denom1
    new["stressp_2"] = (S["stressp_2"][W] + c1nw * (divunw - Dnw)) * denom1
    new["stressp_3"] = (S["stressp_3"][W] + c1sw * (divusw - Dsw)) * denom1
    new["stressp_4"] = (S["stressp_4"][W] + c1se * (divuse - Dse)) * denom1
    new["stressm_1"] = (S["stressm_1"][W] + c0ne * tenne) * denom1
    new["stressm_2"] = (S["stressm_2"][W] + c0nw * tennw) * denom1
    new["stressm_3"] = (S["stressm_3"][W] + c0sw * tensw) * denom1
    new["stressm_4"] = (S["stressm_4"][W] + c0se * tense) * denom1
    new["stress12_1"] = (S["stress12_1"][W] + c0ne * shne * 0.5) * denom1
    new["stress12_2"] = (S["stress12_2"][W] + c0nw * shnw * 0.5) * denom1
    new["stress12_3"] = (S["stress12_3"][W] + c0sw * shsw * 0.5) * denom1
    new["stress12_4"] = (S["stress12_4"][W] + c0se * shse * 0.5) * denom1
    for n, val in new.items():
        S[n][W] = np.where(m, val, S[n][W])
    p1, p2, p3, p4 = (S[f"stressp_{c}"][W] for c in (1, 2, 3, 4))
    m1, m2, m3, m4 = (S[f"stressm_{c}"][W] for c in (1, 2, 3, 4))
    q1, q2, q3, q4 = (S[f"stress12_{c}"][W] for c in (1, 2, 3, 4))
    ssigpn, ssigps, ssigpe, ssigpw = p1 + p2, p3 + p4, p1 + p4, p2 + p3
    ssigp1, ssigp2 = (p1 + p3) * p055, (p2 + p4) * p055
    ssigmn, ssigms, ssigme, ssigmw = m1 + m2, m3 + m4, m1 + m4, m2 + m3
    ssigm1, ssigm2 = (m1 + m3) * p055, (m2 + m4) * p055
    ssig12n, ssig12s, ssig12e, ssig12w = q1 + q2, q3 + q4, q1 + q4, q2 + q3
    ssig121, ssig122 = (q1 + q3) * p111, (q2 + q4) * p111
    csigpne = p111 * p1 + ssigp2 + p027 * p3
    csigpnw = p111 * p2 + ssigp1 + p027 * p4
    csigpsw = p111 * p3 + ssigp2 + p027 * p1
    csigpse = p111 * p4 + ssigp1 + p027 * p2
    csigmne = p111 * m1 + ssigm2 + p027 * m3
    csigmnw = p111 * m2 + ssigm1 + p027 * m4
    csigmsw = p111 * m3 + ssigm2 + p027 * m1
    csigmse = p111 * m4 + ssigm1 + p027 * m2
    csig12ne = p222 * q1 + ssig122 + p055 * q3
    csig12nw = p222 * q2 + ssig121 + p055 * q4
    csig12sw = p222 * q3 + ssig122 + p055 * q1
    csig12se = p222 * q4 + ssig121 + p055 * q2
    str12ew = 0.5 * dxt * (p333 * ssig12e + p166 * ssig12w)
    str12we = 0.5 * dxt * (p333 * ssig12w + p166 * ssig12e)
    str12ns = 0.5 * dyt * (p333 * ssig12n + p166 * ssig12s)
    str12sn = 0.5 * dyt * (p333 * ssig12s + p166 * ssig12n)
    s = [None] * 8
    sp = 0.25 * dyt * (p333 * ssigpn + p166 * ssigps)
    sm = 0.25 * dyt * (p333 * ssigmn + p166 * ssigms)
    s[0] = -sp - sm - str12ew + dxhy * (-csigpne + csigmne) + dyhx * csig12ne
    s[1] = sp + sm - str12we + dxhy * (-csigpnw + csigmnw) + dyhx * csig12nw
    sp = 0.25 * dyt * (p333 * ssigps + p166 * ssigpn)
    sm = 0.25 * dyt * (p333 * ssigms + p166 * ssigmn)
    s[2] = -sp - sm + str12ew + dxhy * (-csigpse + csigmse) + dyhx * csig12se
    s[3] = sp + sm + str12we + dxhy * (-csigpsw + csigmsw) + dyhx * csig12sw
    sp = 0.25 * dxt * (p333 * ssigpe + p166 * ssigpw)
    sm = 0.25 * dxt * (p333 * ssigme + p166 * ssigmw)
    s[4] = -sp + sm - str12ns - dyhx * (csigpne + csigmne) + dxhy * csig12ne
    s[5] = sp - sm - str12sn - dyhx * (csigpse + csigmse) + dxhy * csig12se
    sp = 0.25 * dxt * (p333 * ssigpw + p166 * ssigpe)
    sm = 0.25 * dxt * (p333 * ssigmw + p166 * ssigme)
    s[6] = -sp + sm + str12ns - dyhx * (csigpnw + csigmnw) + dxhy * csig12nw
    s[7] = sp - sm + str12sn - dyhx * (csigpsw + csigmsw) + dxhy * csig12sw
    full = np.zeros((8,) + u.shape)
    for q in range(8):
        full[q][W] = np.where(m, s[q], 0.0)
    return full


def evp(case, cosw=1.0, sinw=0.0, tilt_mode=0, strength_masked=0, dt_init=None, revised_evp=False):
    """One call of evp(dt) on a fixtures.Case (single block); updates case.io / case.out in place."""
    dom, g, inp, io, out = case.dom, case.grid.f, case.inp, case.io, case.out
    dt, ndte = case.dt, case.ndte
    tm = g["tmask"][1:-1, 1:-1] > 0
    xmin = min(g["dxt"][1:-1, 1:-1][tm].min(), g["dyt"][1:-1, 1:-1][tm].min())
    k = evp_parameters(case.dt if dt_init is None else dt_init, ndte, revised_evp, xmin)
    brlx = k[2]
    revp = 1.0 if revised_evp else 0.0
    rhoi, rhos, rhow, gravit = fx.RHOI, fx.RHOS, fx.RHOW, fx.GRAVIT
    a_min, m_min = 1.0e-3, 1.0e-2
    P = (slice(1, -1), slice(1, -1))
    tmask, umask = g["tmask"] > 0, g["umask"] > 0
    S = {n: io[n] for n in fx.STRESS_FIELDS}
    u, v = io["uvel"], io["vvel"]

    for n in ("rdg_conv", "rdg_shear", "divu", "shear", "prs_sig"):            # evp :174-182
        out[n][...] = 0.0
    # evp_prep1 (ice_dyn_shared.F90:314-363)
    tmass = np.where(tmask, rhoi * inp["vice"] + rhos * inp["vsno"], 0.0)
    tmphm = tmask & (inp["aice"] > a_min) & (tmass > m_min)
    icetmask = np.zeros(tmask.shape, dtype=np.int32)
    any9 = np.zeros((dom.ny, dom.nx), dtype=bool)
    for dj in (-1, 0, 1):
        for di in (-1, 0, 1):
            any9 |= _sh(tmphm, di, dj)
    icetmask[P] = (any9 & tmask[P]).astype(np.int32)
    im = icetmask.astype("f8")
    fx.halo_update(dom, im, fx.CENTER, fx.SCALAR, 0.0)                       # evp :210
    icetmask = im.astype(np.int32)
    umass = to_ugrid(tmass, g["tarea"], g["uarea"])                           # :218-219
    aiu = to_ugrid(inp["aice_init"], g["tarea"], g["uarea"])
    if inp.get("strax") is not None:          # ACCESS, .not. calc_strair (:220-229): the coupler's U-grid stress
        strairx, strairy = inp["strax"].copy(), inp["stray"].copy()
    else:
        w = inp["strairxT"].copy(); fx.halo_update(dom, w, fx.CENTER, fx.VECTOR, 0.0)   # :240-241
        strairx = to_ugrid(w, g["tarea"], g["uarea"])
        w = inp["strairyT"].copy(); fx.halo_update(dom, w, fx.CENTER, fx.VECTOR, 0.0)
        strairy = to_ugrid(w, g["tarea"], g["uarea"])
    # evp_prep2 (ice_dyn_shared.F90:484-612)
    for n in fx.STRESS_FIELDS:
        if revised_evp:
            S[n][...] = 0.0                                                       # :492-504
        else:
            S[n][icetmask == 0] = 0.0
    tact = icetmask == 1
    tact[0, :] = False; tact[:, 0] = False
    old = io["iceumask"] > 0
    ice = np.zeros_like(old)
    ice[P] = umask[P] & (aiu[P] > a_min) & (umass[P] > m_min)
    new_pts = ice & ~old
    u[new_pts] = inp["uocn"][new_pts]; v[new_pts] = inp["vocn"][new_pts]
    off = np.zeros_like(old); off[P] = ~ice[P]
    for a in (u, v, out["strintx"], out["strinty"], out["strocnx"], out["strocny"]):
        a[off] = 0.0
    io["iceumask"][P] = ice[P].astype(np.int32)
    out["uvel_init"][P] = u[P]; out["vvel_init"][P] = v[P]
    z = lambda: np.zeros(u.shape)
    umassdti, waterx, watery, forcex, forcey = z(), z(), z(), z(), z()
    fm = io["fm"]
    uo, vo = inp["uocn"], inp["vocn"]
    umassdti[ice] = umass[ice] / dt
    fm[ice] = g["fcor"][ice] * umass[ice]
    sg = np.copysign(1.0, fm)
    waterx[ice] = (uo * cosw - vo * sinw * sg)[ice]
    watery[ice] = (vo * cosw + uo * sinw * sg)[ice]
    if tilt_mode == 0:
        tx, ty = -fm * vo, fm * uo
    else:
        tx, ty = -gravit * umass * inp["ss_tltx"], -gravit * umass * inp["ss_tlty"]
    out["strtltx"][ice] = tx[ice]; out["strtlty"][ice] = ty[ice]
    forcex[ice] = (strairx + out["strtltx"])[ice]; forcey[ice] = (strairy + out["strtlty"])[ice]
    out["strairx"][...] = strairx; out["strairy"][...] = strairy
    out["tmass"][...] = tmass
    # ice_strength stand-in (:291) + halos (:311-315)
    strength = np.zeros(u.shape)
    if strength_masked:
        sel = tact.copy(); sel[-1, :] = False; sel[:, -1] = False
        strength[sel] = inp["strength"][sel]
    else:
        strength[P] = inp["strength"][P]
    fx.halo_update(dom, strength, fx.CENTER, fx.SCALAR, 0.0)
    fx.halo_update(dom, u, fx.NECORNER, fx.VECTOR, 0.0, first_pair=2)         # fld2: ice_HaloUpdate3DR8
    fx.halo_update(dom, v, fx.NECORNER, fx.VECTOR, 0.0, first_pair=2)
    out["strength_out"][...] = strength
    out["icetmask"][...] = icetmask
    Cw = inp["Cw"]
    uarear = g["uarear"]
    for ksub in range(1, ndte + 1):                                           # :336-410
        s = stress(k, ksub, ndte, tact, u, v, g, strength, S, out)
        uold, vold = u.copy(), v.copy()
        vrel = aiu * rhow * Cw * np.sqrt((uo - uold) * (uo - uold) + (vo - vold) * (vo - vold))
        taux, tauy = vrel * waterx, vrel * watery
        cca = (brlx + revp) * umassdti + vrel * cosw
        ccb = fm + np.copysign(1.0, fm) * vrel * sinw
        ab2 = cca * cca + ccb * ccb
        sx, sy = np.zeros(u.shape), np.zeros(u.shape)
        sx[P] = uarear[P] * (_sh(s[0], 0, 0) + _sh(s[1], 1, 0) + _sh(s[2], 0, 1) + _sh(s[3], 1, 1))
        sy[P] = uarear[P] * (_sh(s[4], 0, 0) + _sh(s[5], 0, 1) + _sh(s[6], 1, 0) + _sh(s[7], 1, 1))
        cc1 = sx + forcex + taux + umassdti * (brlx * uold + revp * out["uvel_init"])
        cc2 = sy + forcey + tauy + umassdti * (brlx * vold + revp * out["vvel_init"])
        with np.errstate(divide="ignore", invalid="ignore"):
            un = (cca * cc1 + ccb * cc2) / ab2
            vn = (cca * cc2 - ccb * cc1) / ab2
        u[ice] = un[ice]; v[ice] = vn[ice]
        out["strintx"][ice] = sx[ice]; out["strinty"][ice] = sy[ice]
        out["strocnx"][ice] = taux[ice]; out["strocny"][ice] = tauy[ice]
        fx.halo_update(dom, u, fx.NECORNER, fx.VECTOR, 0.0, first_pair=2)
        fx.halo_update(dom, v, fx.NECORNER, fx.VECTOR, 0.0, first_pair=2)
    if dom.grid_tripole:                                                       # :416-481
        def hs(a1, a2):
            top = a2[dom.ny, 1:dom.nx + 1].copy()
            a1[0, :] = 0.0
            ig = np.arange(0, dom.nx + 2); ig[0] = dom.nx; ig[-1] = 1
            a1[dom.ny + 1, :] = top[(dom.nx - ig + 1) - 1]
        for fam in ("stressp", "stressm", "stress12"):
            hs(S[f"{fam}_1"], S[f"{fam}_3"]); hs(S[f"{fam}_3"], S[f"{fam}_1"])
            hs(S[f"{fam}_2"], S[f"{fam}_4"]); hs(S[f"{fam}_4"], S[f"{fam}_2"])
    # evp_finish (:806-842) + u2tgrid_vector
    vrel = rhow * Cw * np.sqrt((uo - u) * (uo - u) + (vo - v) * (vo - v))
    vrel = vrel * aiu
    sg = np.copysign(1.0, fm)
    sx = vrel * ((uo - u) * cosw - (vo - v) * sinw * sg)
    sy = vrel * ((vo - v) * cosw + (uo - u) * sinw * sg)
    out["strocnx"][ice] = sx[ice]; out["strocny"][ice] = sy[ice]
    wx, wy = np.zeros(u.shape), np.zeros(u.shape)
    with np.errstate(divide="ignore", invalid="ignore"):
        wx[ice] = (out["strocnx"] / aiu)[ice]; wy[ice] = (out["strocny"] / aiu)[ice]
    out["strocnxT"][...] = wx; out["strocnyT"][...] = wy
    fx.halo_update(dom, wx, fx.NECORNER, fx.VECTOR, 0.0)
    fx.halo_update(dom, wy, fx.NECORNER, fx.VECTOR, 0.0)
    to_tgrid(wx, g["tarea"], g["uarea"], out["strocnxT"])
    to_tgrid(wy, g["tarea"], g["uarea"], out["strocnyT"])
    return int(tact.sum()), int(ice.sum())
