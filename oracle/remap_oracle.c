/*
 * remap_oracle.c -- CPU restatement of departure_points (TEST INFRASTRUCTURE, see remap_oracle.h).
 *
 * Follows, statement by statement: source/ice_transport_remap.F90:1493-1670 (departure_points, nghost = 1) and the
 * two ice_HaloUpdate calls on dpx, dpy (field_loc_NEcorner, field_type_vector) of horizontal_remap (:574-578).
 * Compile with -ffp-contract=off: the products and sums below are the reference's, in its order.
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>

#include "remap_oracle.h"

static const double c0 = 0.0, p5 = 0.5;

int remap_oracle_departure_points(const evp_oracle_cfg *cfg, double dt, int l_dp_midpt, int no_halo,
                                  const double *uvel, const double *vvel, const double *dxu, const double *dyu,
                                  const double *HTN, const double *HTE, double *dpx, double *dpy,
                                  int *l_stop, int *istop, int *jstop)
{
    if (!cfg || !uvel || !vvel || !dxu || !dyu || !HTN || !HTE || !dpx || !dpy || !l_stop || !istop || !jstop) return 1;
    const int nxb = cfg->nx_global + 2, nyb = cfg->ny_global + 2;
    const int ilo = 2, ihi = nxb - 1, jlo = 2, jhi = nyb - 1;              /* Fortran indices, nghost = 1 */
#define A(f, i, j) f[(size_t)((j) - 1) * nxb + ((i) - 1)]
    for (size_t q = 0; q < (size_t)nxb * nyb; ++q) { dpx[q] = c0; dpy[q] = c0; }        /* :1557-1558 */
    *l_stop = 0; *istop = 0; *jstop = 0;

    for (int j = jlo; j <= jhi; ++j)                                       /* :1560-1575 */
        for (int i = ilo; i <= ihi; ++i) {
            A(dpx, i, j) = -dt * A(uvel, i, j);
            A(dpy, i, j) = -dt * A(vvel, i, j);
            if (A(dpx, i, j) < -A(HTN, i, j) || A(dpx, i, j) > A(HTN, i + 1, j) ||
                A(dpy, i, j) < -A(HTE, i, j) || A(dpy, i, j) > A(HTE, i, j + 1)) {
                *l_stop = 1; *istop = i; *jstop = j;
            }
        }
    if (*l_stop) return 0;                                                 /* :1577-1591: report and return */

    if (l_dp_midpt) {                                                      /* :1593-1668 */
        for (int j = jlo; j <= jhi; ++j)
            for (int i = ilo; i <= ihi; ++i) {
                if (A(uvel, i, j) != c0 || A(vvel, i, j) != c0) {
                    int i2 = i, j2 = j;
                    double mpx, mpy, mpxt = c0, mpyt = c0, ump, vmp;
                    A(dpx, i, j) = A(dpx, i, j) / A(dxu, i, j);
                    A(dpy, i, j) = A(dpy, i, j) / A(dyu, i, j);
                    mpx = p5 * A(dpx, i, j);
                    mpy = p5 * A(dpy, i, j);
                    if (mpx >= c0 && mpy >= c0) { i2 = i + 1; j2 = j + 1; mpxt = mpx - p5; mpyt = mpy - p5; }
                    else if (mpx < c0 && mpy < c0) { i2 = i; j2 = j; mpxt = mpx + p5; mpyt = mpy + p5; }
                    else if (mpx >= c0 && mpy < c0) { i2 = i + 1; j2 = j; mpxt = mpx - p5; mpyt = mpy + p5; }
                    else if (mpx < c0 && mpy >= c0) { i2 = i; j2 = j + 1; mpxt = mpx + p5; mpyt = mpy - p5; }
                    ump = A(uvel, i2 - 1, j2 - 1) * (mpxt - p5) * (mpyt - p5)
                        - A(uvel, i2, j2 - 1) * (mpxt + p5) * (mpyt - p5)
                        + A(uvel, i2, j2) * (mpxt + p5) * (mpyt + p5)
                        - A(uvel, i2 - 1, j2) * (mpxt - p5) * (mpyt + p5);
                    vmp = A(vvel, i2 - 1, j2 - 1) * (mpxt - p5) * (mpyt - p5)
                        - A(vvel, i2, j2 - 1) * (mpxt + p5) * (mpyt - p5)
                        + A(vvel, i2, j2) * (mpxt + p5) * (mpyt + p5)
                        - A(vvel, i2 - 1, j2) * (mpxt - p5) * (mpyt + p5);
                    A(dpx, i, j) = -dt * ump;
                    A(dpy, i, j) = -dt * vmp;
                }
            }
    }
#undef A
    if (!no_halo) {                                                        /* horizontal_remap :574-578 */
        if (evp_oracle_halo_r8(cfg, dpx, 1, 1, 0.0)) return 1;
        if (evp_oracle_halo_r8(cfg, dpy, 1, 1, 0.0)) return 1;
    }
    return 0;
}

/* ---- second stage: make_masks + construct_fields + limited_gradient (:867-1484), nghost = 1 ------------------------
 * Index lists are kept as the reference keeps them, so the cell sets of every loop are the reference's. */
static const double c1 = 1.0, puny = 1.0e-11;

typedef struct { int nxb, nyb, ilo, ihi, jlo, jhi; } blk;
#define B2(f, i, j) (f)[(size_t)((j) - 1) * b->nxb + ((i) - 1)]

/* Fortran's min / max of a list: the FIRST smallest / largest argument (matters for the sign of a zero only) */
static double min2(double a, double x) { return x < a ? x : a; }
static double max2(double a, double x) { return x > a ? x : a; }

/* limited_gradient, :1348-1484 */
static void limited_gradient(const blk *b, const double *phi, const double *phimask, const double *cnx,
                             const double *cny, double *gx, double *gy, int *indxi, int *indxj)
{
    const size_t n = (size_t)b->nxb * b->nyb;
    for (size_t q = 0; q < n; ++q) { gx[q] = c0; gy[q] = c0; }
    int icells = 0;
    for (int j = b->jlo; j <= b->jhi; ++j)                  /* jlo-nghost+1 .. jhi+nghost-1 */
        for (int i = b->ilo; i <= b->ihi; ++i)
            if (B2(phimask, i, j) > puny) { indxi[icells] = i; indxj[icells] = j; ++icells; }
    for (int ij = 0; ij < icells; ++ij) {
        const int i = indxi[ij], j = indxj[ij];
        const double p = B2(phi, i, j);
#define NB(di, dj) (B2(phimask, i + (di), j + (dj)) * B2(phi, i + (di), j + (dj)) + (c1 - B2(phimask, i + (di), j + (dj))) * p)
        const double phi_nw = NB(-1, 1), phi_n = NB(0, 1), phi_ne = NB(1, 1), phi_w = NB(-1, 0), phi_e = NB(1, 0),
                     phi_sw = NB(-1, -1), phi_s = NB(0, -1), phi_se = NB(1, -1);
#undef NB
        const double gxtmp = (phi_e - phi_w) * p5, gytmp = (phi_n - phi_s) * p5;
        double pmn = phi_nw, pmx = phi_nw;
        const double lst[8] = {phi_n, phi_ne, phi_w, p, phi_e, phi_sw, phi_s, phi_se};
        for (int k = 0; k < 8; ++k) { pmn = min2(pmn, lst[k]); pmx = max2(pmx, lst[k]); }
        pmn = pmn - p;
        pmx = pmx - p;
        const double cx = B2(cnx, i, j), cy = B2(cny, i, j);
        double w1 = (p5 - cx) * gxtmp + (p5 - cy) * gytmp;
        double w2 = (p5 - cx) * gxtmp - (p5 + cy) * gytmp;
        const double w3 = -(p5 + cx) * gxtmp - (p5 + cy) * gytmp;
        const double w4 = (p5 - cy) * gytmp - (p5 + cx) * gxtmp;
        const double qmn = min2(min2(min2(w1, w2), w3), w4), qmx = max2(max2(max2(w1, w2), w3), w4);
        if (fabs(qmn) > fabs(pmn)) w1 = max2(c0, pmn / qmn); else w1 = c1;
        if (fabs(qmx) > fabs(pmx)) w2 = max2(c0, pmx / qmx); else w2 = c1;
        w1 = min2(w1, w2);
        B2(gx, i, j) = w1 * gxtmp;
        B2(gy, i, j) = w1 * gytmp;
    }
}

int remap_oracle_construct_fields(const evp_oracle_cfg *cfg, int ncat, int ntrace, const int *tracer_type,
                                  const int *depend, const int *has_dependents, int no_halo, const double *hm,
                                  const double *xav, const double *yav, const double *xxav, const double *yyav,
                                  const double *mm, const double *tm, double *mc, double *mx, double *my, double *tc,
                                  double *tx, double *ty, double *mmask_out, double *tmask_out, int *icells_out)
{
    if (!cfg || ncat < 1 || ntrace < 0 || !hm || !xav || !yav || !xxav || !yyav || !mm || !mc || !mx || !my) return 1;
    if (ntrace > 0 && (!tracer_type || !depend || !has_dependents || !tm || !tc || !tx || !ty)) return 1;
    blk bb = {cfg->nx_global + 2, cfg->ny_global + 2, 2, cfg->nx_global + 1, 2, cfg->ny_global + 1};
    const blk *b = &bb;
    const size_t pe = (size_t)b->nxb * b->nyb;
    const double nan = NAN;
    /* work arrays of horizontal_remap / construct_fields; mtxav, mtyav start as NaN: the reference leaves them
     * undefined where it does not set them, and nothing below may read such a value */
    double *mmask = malloc((size_t)(ncat + 1) * pe * sizeof(double));
    double *tmask = malloc((size_t)ncat * (ntrace ? ntrace : 1) * pe * sizeof(double));
    double *mxav = malloc(pe * sizeof(double)), *myav = malloc(pe * sizeof(double));
    double *mtxav = malloc((size_t)(ntrace ? ntrace : 1) * pe * sizeof(double));
    double *mtyav = malloc((size_t)(ntrace ? ntrace : 1) * pe * sizeof(double));
    int *indxi = malloc((size_t)(ncat + 1) * pe * sizeof(int)), *indxj = malloc((size_t)(ncat + 1) * pe * sizeof(int));
    int *li = malloc(pe * sizeof(int)), *lj = malloc(pe * sizeof(int)), *icells = malloc((size_t)(ncat + 1) * sizeof(int));
    if (!mmask || !tmask || !mxav || !myav || !mtxav || !mtyav || !indxi || !indxj || !li || !lj || !icells) return 2;
#define MM(n) (mm + (size_t)(n) * pe)
#define TM(n, nt) (tm + ((size_t)((n) - 1) * ntrace + ((nt) - 1)) * pe)          /* n = 1..ncat, nt = 1..ntrace */
#define TMASK(n, nt) (tmask + ((size_t)((n) - 1) * ntrace + ((nt) - 1)) * pe)

    /* ---- make_masks, :867-1015 ---- */
    for (size_t q = 0; q < (size_t)(ncat + 1) * pe; ++q) { indxi[q] = 0; indxj[q] = 0; }
    icells[0] = 0;                                                         /* open water: every cell of the block */
    for (int j = 1; j <= b->nyb; ++j)
        for (int i = 1; i <= b->nxb; ++i) {
            if (B2(MM(0), i, j) > puny) {
                B2(mmask, i, j) = c1;
                indxi[icells[0]] = i; indxj[icells[0]] = j; ++icells[0];
            } else
                B2(mmask, i, j) = c0;
        }
    for (int n = 1; n <= ncat; ++n) {
        int *xi = indxi + (size_t)n * pe, *xj = indxj + (size_t)n * pe;
        double *mk = mmask + (size_t)n * pe;
        icells[n] = 0;
        for (int j = 1; j <= b->nyb; ++j)
            for (int i = 1; i <= b->nxb; ++i)
                if (B2(MM(n), i, j) > puny) { xi[icells[n]] = i; xj[icells[n]] = j; ++icells[n]; }
        for (size_t q = 0; q < pe; ++q) mk[q] = c0;
        for (int ij = 0; ij < icells[n]; ++ij) B2(mk, xi[ij], xj[ij]) = c1;
        for (int nt = 1; nt <= ntrace; ++nt) {
            double *tk = TMASK(n, nt);
            for (size_t q = 0; q < pe; ++q) tk[q] = c0;
            if (has_dependents[nt - 1])
                for (int ij = 0; ij < icells[n]; ++ij)
                    if (fabs(B2(TM(n, nt), xi[ij], xj[ij])) > puny) B2(tk, xi[ij], xj[ij]) = c1;
        }
        icells[n] = 0;                                                     /* redo the list: physical cells only */
        for (int j = b->jlo; j <= b->jhi; ++j)
            for (int i = b->ilo; i <= b->ihi; ++i)
                if (B2(MM(n), i, j) > puny) { xi[icells[n]] = i; xj[icells[n]] = j; ++icells[n]; }
    }

    /* ---- construct_fields, :1024-1340, once per category (horizontal_remap :473-511) ---- */
    for (int n = 0; n <= ncat; ++n) {
        const int *xi = indxi + (size_t)n * pe, *xj = indxj + (size_t)n * pe;
        const double *m = MM(n), *mk = mmask + (size_t)n * pe;
        double *c = mc + (size_t)n * pe, *gx = mx + (size_t)n * pe, *gy = my + (size_t)n * pe;
        const int with_tracers = n >= 1 && ntrace > 0;                     /* present(tm) */
        for (size_t q = 0; q < pe; ++q) { c[q] = c0; gx[q] = c0; gy[q] = c0; mxav[q] = c0; myav[q] = c0; }
        if (with_tracers)
            for (int nt = 1; nt <= ntrace; ++nt) {
                const size_t o = ((size_t)(n - 1) * ntrace + (nt - 1)) * pe;
                for (size_t q = 0; q < pe; ++q) { tc[o + q] = c0; tx[o + q] = c0; ty[o + q] = c0; }
            }
        limited_gradient(b, m, hm, xav, yav, gx, gy, li, lj);
        for (int ij = 0; ij < icells[n]; ++ij) B2(c, xi[ij], xj[ij]) = B2(m, xi[ij], xj[ij]);
        if (!with_tracers) continue;
        for (size_t q = 0; q < (size_t)ntrace * pe; ++q) { mtxav[q] = nan; mtyav[q] = nan; }
        for (int ij = 0; ij < icells[n]; ++ij) {
            const int i = xi[ij], j = xj[ij];
            B2(mxav, i, j) = (B2(gx, i, j) * B2(xxav, i, j) + B2(c, i, j) * B2(xav, i, j)) / B2(m, i, j);
            B2(myav, i, j) = (B2(gy, i, j) * B2(yyav, i, j) + B2(c, i, j) * B2(yav, i, j)) / B2(m, i, j);
        }
        for (int nt = 1; nt <= ntrace; ++nt) {
            const size_t o = ((size_t)(n - 1) * ntrace + (nt - 1)) * pe;
            const double *t = TM(n, nt);
            double *tcn = tc + o, *txn = tx + o, *tyn = ty + o;
            double *ax = mtxav + (size_t)(nt - 1) * pe, *ay = mtyav + (size_t)(nt - 1) * pe;
            if (tracer_type[nt - 1] == 1) {
                limited_gradient(b, t, mk, mxav, myav, txn, tyn, li, lj);
                if (has_dependents[nt - 1]) {
                    const double *tk = TMASK(n, nt);
                    for (size_t q = 0; q < pe; ++q) { ax[q] = c0; ay[q] = c0; }
                    for (int ij = 0; ij < icells[n]; ++ij) {
                        const int i = xi[ij], j = xj[ij];
                        B2(tcn, i, j) = B2(t, i, j) - B2(txn, i, j) * B2(mxav, i, j) - B2(tyn, i, j) * B2(myav, i, j);
                        if (B2(tk, i, j) > puny) {
                            const double w1 = B2(c, i, j) * B2(tcn, i, j);
                            const double w2 = B2(c, i, j) * B2(txn, i, j) + B2(gx, i, j) * B2(tcn, i, j);
                            const double w3 = B2(c, i, j) * B2(tyn, i, j) + B2(gy, i, j) * B2(tcn, i, j);
                            const double w7 = c1 / (B2(m, i, j) * B2(t, i, j));
                            B2(ax, i, j) = (w1 * B2(xav, i, j) + w2 * B2(xxav, i, j)) * w7;
                            B2(ay, i, j) = (w1 * B2(yav, i, j) + w3 * B2(yyav, i, j)) * w7;
                        }
                    }
                } else {
                    for (int ij = 0; ij < icells[n]; ++ij) {
                        const int i = xi[ij], j = xj[ij];
                        B2(tcn, i, j) = B2(t, i, j) - B2(txn, i, j) * B2(mxav, i, j) - B2(tyn, i, j) * B2(myav, i, j);
                    }
                }
            } else if (tracer_type[nt - 1] == 2) {
                const int nt1 = depend[nt - 1];
                const double *ax1 = mtxav + (size_t)(nt1 - 1) * pe, *ay1 = mtyav + (size_t)(nt1 - 1) * pe;
                limited_gradient(b, t, TMASK(n, nt1), ax1, ay1, txn, tyn, li, lj);
                for (int ij = 0; ij < icells[n]; ++ij) {
                    const int i = xi[ij], j = xj[ij];
                    B2(tcn, i, j) = B2(t, i, j) - B2(txn, i, j) * B2(ax1, i, j) - B2(tyn, i, j) * B2(ay1, i, j);
                }
            } else if (tracer_type[nt - 1] == 3) {
                for (int ij = 0; ij < icells[n]; ++ij) B2(tcn, xi[ij], xj[ij]) = B2(t, xi[ij], xj[ij]);
            }
        }
    }
    if (mmask_out) for (size_t q = 0; q < (size_t)(ncat + 1) * pe; ++q) mmask_out[q] = mmask[q];
    if (tmask_out) for (size_t q = 0; q < (size_t)ncat * ntrace * pe; ++q) tmask_out[q] = tmask[q];
    if (icells_out) for (int n = 0; n <= ncat; ++n) icells_out[n] = icells[n];
    free(mmask); free(tmask); free(mxav); free(myav); free(mtxav); free(mtyav); free(indxi); free(indxj); free(li);
    free(lj); free(icells);
#undef MM
#undef TM
#undef TMASK

    if (!no_halo) {                                                        /* horizontal_remap :579-584, :616-621 */
        for (int n = 0; n <= ncat; ++n) {
            if (evp_oracle_halo_r8(cfg, mc + (size_t)n * pe, 0, 0, 0.0)) return 1;
            if (evp_oracle_halo_r8(cfg, mx + (size_t)n * pe, 0, 1, 0.0)) return 1;
            if (evp_oracle_halo_r8(cfg, my + (size_t)n * pe, 0, 1, 0.0)) return 1;
        }
        for (size_t k = 0; k < (size_t)ncat * ntrace; ++k) {
            if (evp_oracle_halo_r8(cfg, tc + k * pe, 0, 0, 0.0)) return 1;
            if (evp_oracle_halo_r8(cfg, tx + k * pe, 0, 1, 0.0)) return 1;
            if (evp_oracle_halo_r8(cfg, ty + k * pe, 0, 1, 0.0)) return 1;
        }
    }
    return 0;
}
