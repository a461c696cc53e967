"""Independent vectorised numpy restatement of ice_strength (TEST INFRASTRUCTURE; written from the
Fortran, not from strength_oracle.c): source/ice_mechred.F90:2111-2269 with ridge_itd :936-1285.
Must agree with the C oracle to the last bit where libm's exp / sqrt agree with numpy's (they are
the same libm here)."""
import numpy as np

PUNY = 1.0e-11
GSTAR, ASTAR, MAXRAFT, HSTAR = 0.15, 0.05, 1.0, 25.0
PSTAR, CSTAR = 2.75e4, 20.0


def ice_strength(aice, vice, aice0, aicen, vicen, cellmask=None, kstrength=1, krdg_partic=1,
                 krdg_redist=1, mu_rdg=3.0, Cf=17.0, rhow=1026.0, rhoi=917.0, gravit=9.80616):
    ncat = aicen.shape[0]
    strength = np.zeros(aice.shape)
    if kstrength != 1:                                                  # :2259-2265
        P = (slice(1, -1), slice(1, -1))
        strength[P] = PSTAR * vice[P] * np.exp(-CSTAR * (1.0 - aice[P]))
        return strength
    Cp = 0.5 * gravit * (rhow - rhoi) * rhoi / rhow                     # :68
    with np.errstate(divide="ignore", invalid="ignore"):
        # Gsum(-1:ncat) -> G[0:ncat+2]
        G = np.empty((ncat + 2,) + aice.shape)
        G[0] = 0.0
        G[1] = np.where(aice0 > PUNY, aice0, G[0])                      # :1030-1038
        for n in range(1, ncat + 1):                                    # :1040-1053
            G[n + 1] = np.where(aicen[n - 1] > PUNY, G[n] + aicen[n - 1], G[n])
        work = 1.0 / G[ncat + 1]                                        # :1058
        for n in range(0, ncat + 1):
            G[n + 1] = G[n + 1] * work                                  # :1062 (Gsum(-1) untouched)
        apartic = np.zeros((ncat + 1,) + aice.shape)
        if krdg_partic == 0:                                            # :1085-1095
            gi = 1.0 / GSTAR
            for n in range(0, ncat + 1):
                lo, hi = G[n], G[n + 1]
                a1 = gi * (hi - lo) * (2.0 - (lo + hi) * gi)
                a2 = gi * (GSTAR - lo) * (2.0 - (lo + GSTAR) * gi)
                apartic[n] = np.where(hi < GSTAR, a1, np.where(lo < GSTAR, a2, 0.0))
        else:                                                           # :1108-1122
            ai = 1.0 / ASTAR
            xtmp = 1.0 / (1.0 - np.exp(-ai))
            E = np.exp(-G * ai) * xtmp
            for n in range(0, ncat + 1):
                apartic[n] = E[n] - E[n + 1]
        hrmin = np.zeros((ncat,) + aice.shape)
        hrmax = np.zeros_like(hrmin)
        hrexp = np.zeros_like(hrmin)
        krdg = np.ones_like(hrmin)
        for n in range(ncat):
            ok = aicen[n] > PUNY
            hi = vicen[n] / aicen[n]
            if krdg_redist == 0:                                        # :1153-1170
                mn = np.minimum(2.0 * hi, hi + MAXRAFT)
                mx = np.maximum(2.0 * np.sqrt(HSTAR * hi), mn + PUNY)
                hrmin[n] = np.where(ok, mn, 0.0)
                hrmax[n] = np.where(ok, mx, 0.0)
                krdg[n] = np.where(ok, (0.5 * (mn + mx)) / hi, 1.0)
            else:                                                       # :1203-1224
                hi = np.maximum(hi, PUNY)
                mn = np.minimum(2.0 * hi, hi + MAXRAFT)
                ex = mu_rdg * np.sqrt(hi)
                hrmin[n] = np.where(ok, mn, 0.0)
                hrexp[n] = np.where(ok, ex, 0.0)
                krdg[n] = np.where(ok, (mn + ex) / hi, 1.0)
        aksum = apartic[0].copy()                                       # :1237-1245
        for n in range(ncat):
            aksum = aksum + apartic[n + 1] * (1.0 - 1.0 / krdg[n])
        s = np.zeros(aice.shape)
        for n in range(ncat):                                           # :2199-2240
            use = (aicen[n] > PUNY) & (apartic[n + 1] > 0.0)
            hi = vicen[n] / aicen[n]
            if krdg_redist == 0:
                h2 = (1.0 / 3.0) * (hrmax[n] * hrmax[n] * hrmax[n] - hrmin[n] * hrmin[n] * hrmin[n]) \
                     / (hrmax[n] - hrmin[n])
            else:
                h2 = hrmin[n] * hrmin[n] + 2.0 * hrmin[n] * hrexp[n] + 2.0 * hrexp[n] * hrexp[n]
            d = -hi * hi + h2 / krdg[n]
            s = np.where(use, s + apartic[n + 1] * d, s)
        res = Cf * Cp * s / aksum                                       # :2249
    if cellmask is None:
        return res
    return np.where(cellmask != 0, res, 0.0)
