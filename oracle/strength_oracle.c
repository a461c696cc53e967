/*
 * strength_oracle.c -- CPU restatement of CICE5's ice_strength (TEST INFRASTRUCTURE, see
 * evp_oracle.h: only tests/, smoke() and bench.py's CPU legs may load it).
 *
 * Follows, statement by statement (citations relative to the reference checkout):
 *   source/ice_mechred.F90:2111-2269  ice_strength   (kstrength = 1 Rothrock 75, = 0 Hibler 79)
 *   source/ice_mechred.F90:758-812    asum_ridging   (computed by the reference, not used by the
 *                                                     strength itself; left out)
 *   source/ice_mechred.F90:936-1285   ridge_itd      (krdg_partic 0 / 1, krdg_redist 0 / 1)
 *   source/ice_mechred.F90:65-82      constants Cp, Gstar, astar, maxraft, Hstar, Pstar, Cstar
 * call site in evp: source/ice_dyn_evp.F90:291-301 -- on the icellt list (cells with
 * icetmask == 1, physical cells and the N/E ghost ring), strength = 0 everywhere else
 * (ice_mechred.F90:2175); evp then fills the ghosts with a halo update (:311).
 *
 * Pinned bit for bit to the reference's own ice_strength / ridge_itd text, executed mechanically
 * (tests/golden/make_reference_vectors.py -> tests/test_oracle_reference_text.py, all four
 * krdg_partic x krdg_redist combinations and kstrength = 0); an independent numpy restatement is
 * oracle/strength_numpy.py.  sqrt() is IEEE; exp() is libm's (cfg->exp_mode = 0: what the mechanical
 * execution of the reference text used) or evp_exp (exp_mode = 1: correctly rounded, the same definition the
 * device kernel compiles -- the two differ where glibc's exp is not correctly rounded, about 1 argument in
 * 1000, by one ulp).
 *
 * Arrays: (ny+2) x (nx+2) global-with-ghost planes, i fastest; aicen / vicen are ncat such planes
 * one after the other (Fortran (nx_block, ny_block, ncat)).  `cellmask` selects the cells of the
 * list (non-zero = in the list); NULL = all cells.
 */
#include <math.h>
#include <stddef.h>

#include "strength_oracle.h"
/* the one definition of the correctly rounded exp shared with the device kernels (a header of plain IEEE
 * operations, no device code): exp_mode = 1 */
#include "../cice5_b200/csrc/evp_exp.h"

#define NCAT_MAX 16

static const double c0 = 0.0, c1 = 1.0, c2 = 2.0, p5 = 0.5, p333 = 1.0 / 3.0;
static const double puny = 1.0e-11;

int strength_oracle_run(const strength_oracle_cfg *cfg, const double *aice, const double *vice,
                        const double *aice0, const double *aicen, const double *vicen,
                        const int *cellmask, double *strength)
{
    const int ncat = cfg->ncat;
    const size_t n = (size_t)(cfg->nx + 2) * (size_t)(cfg->ny + 2);
    /* ice_mechred.F90:65-82 */
    const double Cp = p5 * cfg->gravit * (cfg->rhow - cfg->rhoi) * cfg->rhoi / cfg->rhow;
    const double Gstar = 0.15, astar = 0.05, maxraft = c1, Hstar = 25.0;
    const double Pstar = 2.75e4, Cstar = 20.0;
    const double Gstari = c1 / Gstar, astari = c1 / astar;                /* :982-984 */
    const int cr = cfg->exp_mode == 1;
#define EXPF(x) (cr ? evp_exp(x) : exp(x))
    if (ncat < 1 || ncat > NCAT_MAX) return 1;

    for (size_t q = 0; q < n; ++q) strength[q] = c0;                       /* :2175 */

    if (cfg->kstrength != 1) {                                             /* Hibler 79, :2259-2265 */
        const int gx = cfg->nx + 2;
        for (int j = 1; j <= cfg->ny; ++j)                                 /* jlo..jhi, ilo..ihi */
            for (int i = 1; i <= cfg->nx; ++i) {
                const size_t q = (size_t)j * gx + i;
                strength[q] = Pstar * vice[q] * EXPF(-Cstar * (c1 - aice[q]));
            }
        return 0;
    }

#pragma omp parallel for schedule(static)
    for (long long qq = 0; qq < (long long)n; ++qq) {
        const size_t q = (size_t)qq;
        double GsumA[NCAT_MAX + 2], *Gsum = GsumA + 1;                     /* Gsum(-1:ncat) */
        double apartic[NCAT_MAX + 1], hrmin[NCAT_MAX + 1], hrmax[NCAT_MAX + 1];
        double hrexp[NCAT_MAX + 1], krdg[NCAT_MAX + 1];
        double aksum, work, str;
        if (cellmask && !cellmask[q]) continue;

        /* ---- ridge_itd, :1000-1018 ---- */
        Gsum[-1] = c0;
        Gsum[0] = c1;
        apartic[0] = c0;
        for (int k = 1; k <= ncat; ++k) {
            Gsum[k] = c1; apartic[k] = c0; hrmin[k] = c0; hrmax[k] = c0; hrexp[k] = c0; krdg[k] = c1;
        }
        /* cumulative thickness distribution, :1030-1053 */
        if (aice0[q] > puny) Gsum[0] = aice0[q];
        else Gsum[0] = Gsum[-1];
        for (int k = 1; k <= ncat; ++k) {
            const double a = aicen[(size_t)(k - 1) * n + q];
            if (a > puny) Gsum[k] = Gsum[k - 1] + a;
            else Gsum[k] = Gsum[k - 1];
        }
        work = c1 / Gsum[ncat];                                            /* :1058 */
        for (int k = 0; k <= ncat; ++k) Gsum[k] = Gsum[k] * work;          /* :1062 */

        if (cfg->krdg_partic == 0) {                                       /* Thorndike 75, :1085-1095 */
            for (int k = 0; k <= ncat; ++k) {
                if (Gsum[k] < Gstar)
                    apartic[k] = Gstari * (Gsum[k] - Gsum[k - 1]) * (c2 - (Gsum[k - 1] + Gsum[k]) * Gstari);
                else if (Gsum[k - 1] < Gstar)
                    apartic[k] = Gstari * (Gstar - Gsum[k - 1]) * (c2 - (Gsum[k - 1] + Gstar) * Gstari);
            }
        } else {                                                           /* exponential, :1108-1122 */
            const double xtmp = c1 / (c1 - EXPF(-astari));
            for (int k = -1; k <= ncat; ++k) Gsum[k] = EXPF(-Gsum[k] * astari) * xtmp;
            for (int k = 0; k <= ncat; ++k) apartic[k] = Gsum[k - 1] - Gsum[k];
        }

        if (cfg->krdg_redist == 0) {                                       /* Hibler 80, :1153-1170 */
            for (int k = 1; k <= ncat; ++k) {
                const double a = aicen[(size_t)(k - 1) * n + q];
                if (a > puny) {
                    const double hi = vicen[(size_t)(k - 1) * n + q] / a;
                    double hrmean;
                    hrmin[k] = fmin(c2 * hi, hi + maxraft);
                    hrmax[k] = c2 * sqrt(Hstar * hi);
                    hrmax[k] = fmax(hrmax[k], hrmin[k] + puny);
                    hrmean = p5 * (hrmin[k] + hrmax[k]);
                    krdg[k] = hrmean / hi;
                }
            }
        } else {                                                           /* exponential, :1203-1224 */
            for (int k = 1; k <= ncat; ++k) {
                const double a = aicen[(size_t)(k - 1) * n + q];
                if (a > puny) {
                    double hi = vicen[(size_t)(k - 1) * n + q] / a;
                    hi = fmax(hi, puny);
                    hrmin[k] = fmin(c2 * hi, hi + maxraft);
                    hrexp[k] = cfg->mu_rdg * sqrt(hi);
                    krdg[k] = (hrmin[k] + hrexp[k]) / hi;
                }
            }
        }

        aksum = apartic[0];                                                /* :1237 */
        for (int k = 1; k <= ncat; ++k)                                    /* :1243-1244 */
            aksum = aksum + apartic[k] * (c1 - c1 / krdg[k]);

        /* ---- ice_strength, :2199-2253 ---- */
        str = c0;
        for (int k = 1; k <= ncat; ++k) {
            const double a = aicen[(size_t)(k - 1) * n + q];
            if (a > puny && apartic[k] > c0) {
                const double hi = vicen[(size_t)(k - 1) * n + q] / a;
                double h2rdg, dh2rdg;
                if (cfg->krdg_redist == 0)                                 /* :2209-2211 */
                    h2rdg = p333 * (hrmax[k] * hrmax[k] * hrmax[k] - hrmin[k] * hrmin[k] * hrmin[k]) /
                            (hrmax[k] - hrmin[k]);
                else                                                       /* :2229-2231 */
                    h2rdg = hrmin[k] * hrmin[k] + c2 * hrmin[k] * hrexp[k] + c2 * hrexp[k] * hrexp[k];
                dh2rdg = -hi * hi + h2rdg / krdg[k];
                str = str + apartic[k] * dh2rdg;
            }
        }
        strength[q] = cfg->Cf * Cp * str / aksum;                          /* :2249 */
    }
    return 0;
}
#undef EXPF

/* evp_exp on n arguments: lets tests/test_oracle_cpu.py check the shared definition against exact arithmetic */
void strength_oracle_cr_exp(long long n, const double *x, double *y)
{
    for (long long i = 0; i < n; ++i) y[i] = evp_exp(x[i]);
}
