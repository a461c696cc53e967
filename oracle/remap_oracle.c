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

/* ---- third stage: locate_triangles (:1680-3047), triangle_coordinates (:3078-3187), transport_integrals (:3199-3509),
 * update_fields (:3517-3729) -- with the reference's compile-time settings l_fixed_area = .false., integral_order = 3
 * (source/ice_transport_driver.F90:31,47); the l_fixed_area branches are not restated ---------------------------------- */
#define NGROUPS 6
#define NVERT 3
static const double eps16 = 1.0e-16, p4 = 0.4, p6 = 0.6;

/* arrays: xp, yp [ng][nv 0..3][nyb][nxb]; iflux, jflux, triarea [ng][nyb][nxb]; indxi, indxj [ng][nxb*nyb] */
typedef struct {
    double *xp, *yp, *triarea;
    int *iflux, *jflux, *indxi, *indxj;
    int icells[NGROUPS];
} triangles;
#define XP(t, i, j, nv, ng) (t)->xp[(((size_t)((ng) - 1) * (NVERT + 1) + (nv)) * b->nyb + ((j) - 1)) * b->nxb + ((i) - 1)]
#define YP(t, i, j, nv, ng) (t)->yp[(((size_t)((ng) - 1) * (NVERT + 1) + (nv)) * b->nyb + ((j) - 1)) * b->nxb + ((i) - 1)]
#define G3(f, i, j, ng) (f)[((size_t)((ng) - 1) * b->nyb + ((j) - 1)) * b->nxb + ((i) - 1)]

static int locate_triangles(const blk *b, int north, const double *dpx, const double *dpy, const double *dxu,
                            const double *dyu, triangles *t, double *edgearea)
{
    const size_t pe = (size_t)b->nxb * b->nyb;
    double *dx = calloc(pe, sizeof(double)), *dy = calloc(pe, sizeof(double));
    double *fac_c = calloc(pe, sizeof(double)), *fac_l = calloc(pe, sizeof(double)), *fac_r = calloc(pe, sizeof(double));
    double *areafact = calloc((size_t)NGROUPS * pe, sizeof(double)), *areasum = calloc(pe, sizeof(double));
    int *indxid = malloc(pe * sizeof(int)), *indxjd = malloc(pe * sizeof(int));
    if (!dx || !dy || !fac_c || !fac_l || !fac_r || !areafact || !areasum || !indxid || !indxjd) return 2;
    for (int ng = 1; ng <= NGROUPS; ++ng) {                               /* :1800-1818 */
        for (int j = 1; j <= b->nyb; ++j)
            for (int i = 1; i <= b->nxb; ++i) {
                G3(t->triarea, i, j, ng) = c0;
                G3(t->iflux, i, j, ng) = i;
                G3(t->jflux, i, j, ng) = j;
                for (int nv = 0; nv <= NVERT; ++nv) { XP(t, i, j, nv, ng) = c0; YP(t, i, j, nv, ng) = c0; }
            }
    }
    int ib, ie, jb, je, is_tl, js_tl, is_bl, js_bl, is_tr, js_tr, is_br, js_br, is_tc, js_tc, is_bc, js_bc;
    if (north) {                                                          /* :1820-1858 */
        ib = b->ilo; ie = b->ihi; jb = b->jlo - 1; je = b->jhi;
        is_tl = -1; js_tl = 1; is_bl = -1; js_bl = 0; is_tr = 1; js_tr = 1; is_br = 1; js_br = 0;
        is_tc = 0; js_tc = 1; is_bc = 0; js_bc = 0;
        for (int j = jb; j <= je; ++j)
            for (int i = ib; i <= ie; ++i) {
                B2(fac_l, i, j) = B2(dxu, i - 1, j) * B2(dyu, i - 1, j);
                B2(fac_r, i, j) = B2(dxu, i, j) * B2(dyu, i, j);
                B2(fac_c, i, j) = p5 * (B2(fac_l, i, j) + B2(fac_r, i, j));
            }
    } else {                                                              /* :1860-1898 */
        ib = b->ilo - 1; ie = b->ihi; jb = b->jlo; je = b->jhi;
        is_tl = 1; js_tl = 1; is_bl = 0; js_bl = 1; is_tr = 1; js_tr = -1; is_br = 0; js_br = -1;
        is_tc = 1; js_tc = 0; is_bc = 0; js_bc = 0;
        for (int j = jb; j <= je; ++j)
            for (int i = ib; i <= ie; ++i) {
                B2(fac_l, i, j) = B2(dxu, i, j) * B2(dyu, i, j);
                B2(fac_r, i, j) = B2(dxu, i, j - 1) * B2(dyu, i, j - 1);
                B2(fac_c, i, j) = p5 * (B2(fac_l, i, j) + B2(fac_r, i, j));
            }
    }
    int icellsd = 0;                                                      /* :1956-1980: edges with a moving corner */
    for (int j = jb; j <= je; ++j)
        for (int i = ib; i <= ie; ++i) {
            const int il = north ? i - 1 : i, jl = north ? j : j - 1;
            if (B2(dpx, il, jl) != c0 || B2(dpy, il, jl) != c0 || B2(dpx, i, j) != c0 || B2(dpy, i, j) != c0) {
                indxid[icellsd] = i; indxjd[icellsd] = j; ++icellsd;
            }
        }
    for (int j = 1; j <= je; ++j)                                         /* :1986-1991 */
        for (int i = 1; i <= ie; ++i) {
            B2(dx, i, j) = B2(dpx, i, j) / B2(dxu, i, j);
            B2(dy, i, j) = B2(dpy, i, j) / B2(dyu, i, j);
        }
#define TRI(ng, x1, y1, x2, y2, x3, y3, sh, fac) TRI_(ng, x1, y1, x2, y2, x3, y3, sh, fac)
#define TRI_(ng, x1, y1, x2, y2, x3, y3, is, js, fac) do { \
        XP(t, i, j, 1, ng) = (x1); YP(t, i, j, 1, ng) = (y1); XP(t, i, j, 2, ng) = (x2); YP(t, i, j, 2, ng) = (y2); \
        XP(t, i, j, 3, ng) = (x3); YP(t, i, j, 3, ng) = (y3); G3(t->iflux, i, j, ng) = i + (is); \
        G3(t->jflux, i, j, ng) = j + (js); G3(areafact, i, j, ng) = (fac); } while (0)
#define TL is_tl, js_tl
#define BL is_bl, js_bl
#define TR is_tr, js_tr
#define BR is_br, js_br
#define TC is_tc, js_tc
#define BC is_bc, js_bc
    for (int ij = 0; ij < icellsd; ++ij) {                                /* :1993-2925 */
        const int i = indxid[ij], j = indxjd[ij];
        const double xcl = -p5, ycl = c0, xcr = p5, ycr = c0;
        double xdl, ydl, xdr, ydr;
        if (north) {
            xdl = xcl + B2(dx, i - 1, j); ydl = ycl + B2(dy, i - 1, j);
            xdr = xcr + B2(dx, i, j);     ydr = ycr + B2(dy, i, j);
        } else {                                                          /* rotate the trajectory by pi/2 */
            xdl = xcl - B2(dy, i, j);     ydl = ycl + B2(dx, i, j);
            xdr = xcr - B2(dy, i, j - 1); ydr = ycr + B2(dx, i, j - 1);
        }
        const double xdm = p5 * (xdr + xdl), ydm = p5 * (ydr + ydl);
        const double xil = xcl, yil = (xcl * (ydm - ydl) + xdm * ydl - xdl * ydm) / (xdm - xdl);
        const double xir = xcr, yir = (xcr * (ydr - ydm) - xdm * ydr + xdr * ydm) / (xdr - xdm);
        const double md = (ydr - ydl) / (xdr - xdl);
        const double xic = fabs(md) > puny ? xdl - ydl / md : c0, yic = c0;
        const double xicl = xic, yicl = yic, xicr = xic, yicr = yic;     /* l_fixed_area = F */
        const double fl = B2(fac_l, i, j), fr = B2(fac_r, i, j), fc = B2(fac_c, i, j);
        /* TL and BL triangles, :2034-2108 */
        if (yil > c0 && xdl < xcl && ydl >= c0)
            TRI(1, xcl, ycl, xil, yil, xdl, ydl, TL, -fl);
        else if (yil < c0 && xdl < xcl && ydl < c0)
            TRI(1, xcl, ycl, xdl, ydl, xil, yil, BL, fl);
        else if (yil < c0 && xdl < xcl && ydl >= c0) {
            TRI(1, xcl, ycl, xdl, ydl, xic, yic, TL, fl);
            TRI(3, xcl, ycl, xic, yic, xil, yil, BL, fl);
        } else if (yil > c0 && xdl < xcl && ydl < c0) {
            TRI(3, xcl, ycl, xil, yil, xic, yic, TL, -fl);
            TRI(1, xcl, ycl, xic, yic, xdl, ydl, BL, -fl);
        }
        /* TR and BR triangles, :2114-2190 */
        if (yir > c0 && xdr >= xcr && ydr >= c0)
            TRI(2, xcr, ycr, xdr, ydr, xir, yir, TR, -fr);
        else if (yir < c0 && xdr >= xcr && ydr < c0)
            TRI(2, xcr, ycr, xir, yir, xdr, ydr, BR, fr);
        else if (yir < c0 && xdr >= xcr && ydr >= c0) {
            TRI(2, xcr, ycr, xic, yic, xdr, ydr, TR, fr);
            TRI(3, xcr, ycr, xir, yir, xic, yic, BR, fr);
        } else if (yir > c0 && xdr >= xcr && ydr < c0) {
            TRI(3, xcr, ycr, xic, yic, xir, yir, TR, -fr);
            TRI(2, xcr, ycr, xdr, ydr, xic, yic, BR, -fr);
        }
        /* redefine departure points that lie in side cells, :2238-2255 */
        if (xdl < xcl) { xdl = xil; ydl = yil; }
        if (xdr > xcr) { xdr = xir; ydr = yir; }
        /* TC and BC triangles, :2394-2923 */
        if (ydl >= c0 && ydr >= c0 && ydm >= c0) {
            TRI(4, xcl, ycl, xcr, ycr, xdl, ydl, TC, -fc);
            TRI(5, xcr, ycr, xdr, ydr, xdl, ydl, TC, -fc);
            TRI(6, xdl, ydl, xdr, ydr, xdm, ydm, TC, -fc);
        } else if (ydl >= c0 && ydr >= c0 && ydm < c0) {
            TRI(4, xcl, ycl, xicl, yicl, xdl, ydl, TC, -fc);
            TRI(5, xcr, ycr, xdr, ydr, xicr, yicr, TC, -fc);
            TRI(6, xicr, yicr, xicl, yicl, xdm, ydm, BC, fc);
        } else if (ydl < c0 && ydr < c0 && ydm < c0) {
            TRI(4, xcl, ycl, xdl, ydl, xcr, ycr, BC, fc);
            TRI(5, xcr, ycr, xdl, ydl, xdr, ydr, BC, fc);
            TRI(6, xdl, ydl, xdm, ydm, xdr, ydr, BC, fc);
        } else if (ydl < c0 && ydr < c0 && ydm >= c0) {
            TRI(4, xcl, ycl, xdl, ydl, xicl, yicl, BC, fc);
            TRI(5, xcr, ycr, xicr, yicr, xdr, ydr, BC, fc);
            TRI(6, xicl, yicl, xicr, yicr, xdm, ydm, TC, -fc);
        } else if (ydl >= c0 && ydr < c0 && xic >= c0 && ydm >= c0) {
            TRI(4, xcl, ycl, xicr, yicr, xdl, ydl, TC, -fc);
            TRI(5, xcr, ycr, xicr, yicr, xdr, ydr, BC, fr);
            TRI(6, xdl, ydl, xicr, yicr, xdm, ydm, TC, -fc);
        } else if (ydl >= c0 && ydr < c0 && xic >= c0 && ydm < c0) {
            TRI(4, xcl, ycl, xicl, yicl, xdl, ydl, TC, -fc);
            TRI(5, xcr, ycr, xicr, yicr, xdr, ydr, BC, fr);
            TRI(6, xicr, yicr, xicl, yicl, xdm, ydm, BC, fc);
        } else if (ydl >= c0 && ydr < c0 && xic < c0 && ydm < c0) {
            TRI(4, xcl, ycl, xicl, yicl, xdl, ydl, TC, -fl);
            TRI(5, xcr, ycr, xicl, yicl, xdr, ydr, BC, fc);
            TRI(6, xdr, ydr, xicl, yicl, xdm, ydm, BC, fc);
        } else if (ydl >= c0 && ydr < c0 && xic < c0 && ydm >= c0) {
            TRI(4, xcl, ycl, xicl, yicl, xdl, ydl, TC, -fl);
            TRI(5, xcr, ycr, xicr, yicr, xdr, ydr, BC, fc);
            TRI(6, xicl, yicl, xicr, yicr, xdm, ydm, TC, -fc);
        } else if (ydl < c0 && ydr >= c0 && xic < c0 && ydm >= c0) {
            TRI(4, xcl, ycl, xdl, ydl, xicl, yicl, BC, fl);
            TRI(5, xcr, ycr, xdr, ydr, xicl, yicl, TC, -fc);
            TRI(6, xicl, yicl, xdr, ydr, xdm, ydm, TC, -fc);
        } else if (ydl < c0 && ydr >= c0 && xic < c0 && ydm < c0) {
            TRI(4, xcl, ycl, xdl, ydl, xicl, yicl, BC, fl);
            TRI(5, xcr, ycr, xdr, ydr, xicr, yicr, TC, -fc);
            TRI(6, xicr, yicr, xicl, yicl, xdm, ydm, BC, fc);
        } else if (ydl < c0 && ydr >= c0 && xic >= c0 && ydm < c0) {
            TRI(4, xcl, ycl, xdl, ydl, xicr, yicr, BC, fc);
            TRI(5, xcr, ycr, xdr, ydr, xicr, yicr, TC, -fr);
            TRI(6, xicr, yicr, xdl, ydl, xdm, ydm, BC, fc);
        } else if (ydl < c0 && ydr >= c0 && xic >= c0 && ydm >= c0) {
            TRI(4, xcl, ycl, xdl, ydl, xicl, yicl, BC, fc);
            TRI(5, xcr, ycr, xdr, ydr, xicr, yicr, TC, -fr);
            TRI(6, xicl, yicl, xicr, yicr, xdm, ydm, TC, -fc);
        }
    }
#undef TRI
#undef TRI_
#undef TL
#undef BL
#undef TR
#undef BR
#undef TC
#undef BC
    for (int ng = 1; ng <= NGROUPS; ++ng) {                               /* :2929-2952 */
        t->icells[ng - 1] = 0;
        for (int ij = 0; ij < icellsd; ++ij) {
            const int i = indxid[ij], j = indxjd[ij];
            G3(t->triarea, i, j, ng) = p5 * ((XP(t, i, j, 2, ng) - XP(t, i, j, 1, ng)) * (YP(t, i, j, 3, ng) - YP(t, i, j, 1, ng)) -
                                             (YP(t, i, j, 2, ng) - YP(t, i, j, 1, ng)) * (XP(t, i, j, 3, ng) - XP(t, i, j, 1, ng))) *
                                       G3(areafact, i, j, ng);
            if (fabs(G3(t->triarea, i, j, ng)) < eps16 * B2(fac_c, i, j))
                G3(t->triarea, i, j, ng) = c0;
            else {
                const size_t ic = (size_t)(ng - 1) * pe + t->icells[ng - 1]++;
                t->indxi[ic] = i; t->indxj[ic] = j;
            }
            B2(areasum, i, j) = B2(areasum, i, j) + G3(t->triarea, i, j, ng);
        }
    }
    if (edgearea)                                                         /* :2978-2982 */
        for (int ij = 0; ij < icellsd; ++ij) B2(edgearea, indxid[ij], indxjd[ij]) = B2(areasum, indxid[ij], indxjd[ij]);
    for (int ng = 1; ng <= NGROUPS; ++ng)                                 /* into the source cell's coordinates, :2989-3021 */
        for (int nv = 1; nv <= NVERT; ++nv)
            for (int ij = 0; ij < t->icells[ng - 1]; ++ij) {
                const int i = t->indxi[(size_t)(ng - 1) * pe + ij], j = t->indxj[(size_t)(ng - 1) * pe + ij];
                const int ishift = G3(t->iflux, i, j, ng) - i, jshift = G3(t->jflux, i, j, ng) - j;
                if (north) {
                    XP(t, i, j, nv, ng) = XP(t, i, j, nv, ng) - c1 * ishift;
                    YP(t, i, j, nv, ng) = YP(t, i, j, nv, ng) + p5 - c1 * jshift;
                } else {
                    const double w1 = XP(t, i, j, nv, ng);
                    XP(t, i, j, nv, ng) = YP(t, i, j, nv, ng) + p5 - c1 * ishift;
                    YP(t, i, j, nv, ng) = -w1 - c1 * jshift;
                }
            }
    free(dx); free(dy); free(fac_c); free(fac_l); free(fac_r); free(areafact); free(areasum); free(indxid); free(indxjd);
    return 0;
}

/* triangle_coordinates, integral_order = 3 (:3160-3183) */
static void triangle_coordinates(const blk *b, triangles *t)
{
    const size_t pe = (size_t)b->nxb * b->nyb;
    const double p333 = c1 / 3.0;
    for (int ng = 1; ng <= NGROUPS; ++ng)
        for (int ij = 0; ij < t->icells[ng - 1]; ++ij) {
            const int i = t->indxi[(size_t)(ng - 1) * pe + ij], j = t->indxj[(size_t)(ng - 1) * pe + ij];
            XP(t, i, j, 0, ng) = p333 * (XP(t, i, j, 1, ng) + XP(t, i, j, 2, ng) + XP(t, i, j, 3, ng));
            YP(t, i, j, 0, ng) = p333 * (YP(t, i, j, 1, ng) + YP(t, i, j, 2, ng) + YP(t, i, j, 3, ng));
            for (int nv = 1; nv <= NVERT; ++nv) {
                XP(t, i, j, nv, ng) = p4 * XP(t, i, j, nv, ng) + p6 * XP(t, i, j, 0, ng);
                YP(t, i, j, nv, ng) = p4 * YP(t, i, j, nv, ng) + p6 * YP(t, i, j, 0, ng);
            }
        }
}

/* transport_integrals, integral_order = 3 (:3380-3505).  tc == NULL: mass only (open water) */
static int transport_integrals(const blk *b, int ntrace, const int *tracer_type, const int *depend, const triangles *t,
                               const double *mc, const double *mx, const double *my, double *mflx, const double *tc,
                               const double *tx, const double *ty, double *mtflx)
{
    const size_t pe = (size_t)b->nxb * b->nyb;
    const double p5625m = -9.0 / 16.0, p52083 = 25.0 / 48.0, nan = NAN;
    double *w = malloc((6 + 3 * (size_t)(ntrace ? ntrace : 1)) * pe * sizeof(double));
    if (!w) return 2;
    double *msum = w, *mxsum = w + pe, *mysum = w + 2 * pe, *mxxsum = w + 3 * pe, *mxysum = w + 4 * pe, *myysum = w + 5 * pe;
    double *mtsum = w + 6 * pe, *mtxsum = mtsum + (size_t)ntrace * pe, *mtysum = mtxsum + (size_t)ntrace * pe;
    for (size_t q = 0; q < (6 + 3 * (size_t)(ntrace ? ntrace : 1)) * pe; ++q) w[q] = nan;     /* undefined until set */
    for (size_t q = 0; q < pe; ++q) mflx[q] = c0;
    if (tc) for (size_t q = 0; q < (size_t)ntrace * pe; ++q) mtflx[q] = c0;
#define T3(f, i, j, nt) (f)[((size_t)((nt) - 1) * b->nyb + ((j) - 1)) * b->nxb + ((i) - 1)]
    for (int ng = 1; ng <= NGROUPS; ++ng) {
        const int *xi = t->indxi + (size_t)(ng - 1) * pe, *xj = t->indxj + (size_t)(ng - 1) * pe;
        const int n = t->icells[ng - 1];
        for (int ij = 0; ij < n; ++ij) {
            const int i = xi[ij], j = xj[ij], i2 = G3(t->iflux, i, j, ng), j2 = G3(t->jflux, i, j, ng);
            const double x0 = XP(t, i, j, 0, ng), x1 = XP(t, i, j, 1, ng), x2 = XP(t, i, j, 2, ng), x3 = XP(t, i, j, 3, ng);
            const double y0 = YP(t, i, j, 0, ng), y1 = YP(t, i, j, 1, ng), y2 = YP(t, i, j, 2, ng), y3 = YP(t, i, j, 3, ng);
            const double m0 = p5625m * (B2(mc, i2, j2) + x0 * B2(mx, i2, j2) + y0 * B2(my, i2, j2));
            const double m1 = p52083 * (B2(mc, i2, j2) + x1 * B2(mx, i2, j2) + y1 * B2(my, i2, j2));
            const double m2 = p52083 * (B2(mc, i2, j2) + x2 * B2(mx, i2, j2) + y2 * B2(my, i2, j2));
            const double m3 = p52083 * (B2(mc, i2, j2) + x3 * B2(mx, i2, j2) + y3 * B2(my, i2, j2));
            B2(msum, i, j) = m0 + m1 + m2 + m3;
            B2(mflx, i, j) = B2(mflx, i, j) + G3(t->triarea, i, j, ng) * B2(msum, i, j);
            double w0 = m0 * x0, w1 = m1 * x1, w2 = m2 * x2, w3 = m3 * x3;
            B2(mxsum, i, j) = w0 + w1 + w2 + w3;
            B2(mxxsum, i, j) = w0 * x0 + w1 * x1 + w2 * x2 + w3 * x3;
            B2(mxysum, i, j) = w0 * y0 + w1 * y1 + w2 * y2 + w3 * y3;
            w0 = m0 * y0; w1 = m1 * y1; w2 = m2 * y2; w3 = m3 * y3;
            B2(mysum, i, j) = w0 + w1 + w2 + w3;
            B2(myysum, i, j) = w0 * y0 + w1 * y1 + w2 * y2 + w3 * y3;
        }
        if (!tc) continue;
        for (int nt = 1; nt <= ntrace; ++nt) {
            const int type = tracer_type[nt - 1], nt1 = depend[nt - 1];
            for (int ij = 0; ij < n; ++ij) {
                const int i = xi[ij], j = xj[ij], i2 = G3(t->iflux, i, j, ng), j2 = G3(t->jflux, i, j, ng);
                if (type == 1) {
                    T3(mtsum, i, j, nt) = B2(msum, i, j) * T3(tc, i2, j2, nt) + B2(mxsum, i, j) * T3(tx, i2, j2, nt) +
                                          B2(mysum, i, j) * T3(ty, i2, j2, nt);
                    T3(mtflx, i, j, nt) = T3(mtflx, i, j, nt) + G3(t->triarea, i, j, ng) * T3(mtsum, i, j, nt);
                    T3(mtxsum, i, j, nt) = B2(mxsum, i, j) * T3(tc, i2, j2, nt) + B2(mxxsum, i, j) * T3(tx, i2, j2, nt) +
                                           B2(mxysum, i, j) * T3(ty, i2, j2, nt);
                    T3(mtysum, i, j, nt) = B2(mysum, i, j) * T3(tc, i2, j2, nt) + B2(mxysum, i, j) * T3(tx, i2, j2, nt) +
                                           B2(myysum, i, j) * T3(ty, i2, j2, nt);
                } else if (type == 2) {
                    T3(mtsum, i, j, nt) = T3(mtsum, i, j, nt1) * T3(tc, i2, j2, nt) + T3(mtxsum, i, j, nt1) * T3(tx, i2, j2, nt) +
                                          T3(mtysum, i, j, nt1) * T3(ty, i2, j2, nt);
                    T3(mtflx, i, j, nt) = T3(mtflx, i, j, nt) + G3(t->triarea, i, j, ng) * T3(mtsum, i, j, nt);
                } else if (type == 3) {
                    T3(mtsum, i, j, nt) = T3(mtsum, i, j, nt1) * T3(tc, i2, j2, nt);
                    T3(mtflx, i, j, nt) = T3(mtflx, i, j, nt) + G3(t->triarea, i, j, ng) * T3(mtsum, i, j, nt);
                }
            }
        }
    }
    free(w);
    return 0;
}

/* update_fields (:3517-3729).  tm == NULL: mass only.  *l_stop etc. as the routine leaves them (last offending cell) */
static int update_fields(const blk *b, int ntrace, const int *tracer_type, const int *depend, const double *tarear,
                         int *l_stop, int *istop, int *jstop, const double *mflxe, const double *mflxn, double *mm,
                         const double *mtflxe, const double *mtflxn, double *tm)
{
    const size_t pe = (size_t)b->nxb * b->nyb;
    double *mtold = tm ? malloc((size_t)(ntrace ? ntrace : 1) * pe * sizeof(double)) : NULL;
    int *xi = malloc(pe * sizeof(int)), *xj = malloc(pe * sizeof(int));
    if ((tm && !mtold) || !xi || !xj) return 2;
    if (tm)
        for (int nt = 1; nt <= ntrace; ++nt) {
            const int type = tracer_type[nt - 1], nt1 = depend[nt - 1], nt2 = type == 3 ? depend[nt1 - 1] : 0;
            for (int j = b->jlo; j <= b->jhi; ++j)
                for (int i = b->ilo; i <= b->ihi; ++i) {
                    if (type == 1) T3(mtold, i, j, nt) = B2(mm, i, j) * T3(tm, i, j, nt);
                    else if (type == 2) T3(mtold, i, j, nt) = B2(mm, i, j) * T3(tm, i, j, nt1) * T3(tm, i, j, nt);
                    else if (type == 3) T3(mtold, i, j, nt) = B2(mm, i, j) * T3(tm, i, j, nt2) * T3(tm, i, j, nt1) * T3(tm, i, j, nt);
                }
        }
    for (int j = b->jlo; j <= b->jhi; ++j)
        for (int i = b->ilo; i <= b->ihi; ++i) {
            const double w1 = B2(mflxe, i, j) - B2(mflxe, i - 1, j) + B2(mflxn, i, j) - B2(mflxn, i, j - 1);
            B2(mm, i, j) = B2(mm, i, j) - w1 * B2(tarear, i, j);
            if (B2(mm, i, j) < -puny) { *l_stop = 1; *istop = i; *jstop = j; }
            else if (B2(mm, i, j) < c0) B2(mm, i, j) = c0;
        }
    if (*l_stop) { free(mtold); free(xi); free(xj); return 0; }
    if (tm) {
        int icells = 0;
        for (int j = b->jlo; j <= b->jhi; ++j)
            for (int i = b->ilo; i <= b->ihi; ++i)
                if (B2(mm, i, j) > c0) { xi[icells] = i; xj[icells] = j; ++icells; }
        for (int nt = 1; nt <= ntrace; ++nt) {
            const int type = tracer_type[nt - 1], nt1 = depend[nt - 1], nt2 = type == 3 ? depend[nt1 - 1] : 0;
            for (int j = b->jlo; j <= b->jhi; ++j)
                for (int i = b->ilo; i <= b->ihi; ++i) T3(tm, i, j, nt) = c0;
            for (int ij = 0; ij < icells; ++ij) {
                const int i = xi[ij], j = xj[ij];
                if (type == 1) {
                    const double w1 = T3(mtflxe, i, j, nt) - T3(mtflxe, i - 1, j, nt) + T3(mtflxn, i, j, nt) - T3(mtflxn, i, j - 1, nt);
                    T3(tm, i, j, nt) = (T3(mtold, i, j, nt) - w1 * B2(tarear, i, j)) / B2(mm, i, j);
                } else if (type == 2) {
                    if (fabs(T3(tm, i, j, nt1)) > c0) {
                        const double w1 = T3(mtflxe, i, j, nt) - T3(mtflxe, i - 1, j, nt) + T3(mtflxn, i, j, nt) - T3(mtflxn, i, j - 1, nt);
                        T3(tm, i, j, nt) = (T3(mtold, i, j, nt) - w1 * B2(tarear, i, j)) / (B2(mm, i, j) * T3(tm, i, j, nt1));
                    }
                } else if (type == 3) {
                    if (fabs(T3(tm, i, j, nt1)) > c0 && fabs(T3(tm, i, j, nt2)) > c0) {
                        const double w1 = T3(mtflxe, i, j, nt) - T3(mtflxe, i - 1, j, nt) + T3(mtflxn, i, j, nt) - T3(mtflxn, i, j - 1, nt);
                        T3(tm, i, j, nt) = (T3(mtold, i, j, nt) - w1 * B2(tarear, i, j)) /
                                           (B2(mm, i, j) * T3(tm, i, j, nt2) * T3(tm, i, j, nt1));
                    }
                }
            }
        }
    }
    free(mtold); free(xi); free(xj);
    return 0;
}

/* ---- exported: the routines one by one (pinned to the reference text), then all of horizontal_remap ------------------ */
static void tri_alloc(triangles *t, const blk *b, double *xp, double *yp, double *triarea, int *iflux, int *jflux)
{
    const size_t pe = (size_t)b->nxb * b->nyb;
    t->xp = xp; t->yp = yp; t->triarea = triarea; t->iflux = iflux; t->jflux = jflux;
    t->indxi = malloc((size_t)NGROUPS * pe * sizeof(int));
    t->indxj = malloc((size_t)NGROUPS * pe * sizeof(int));
}
/* the lists of transport_integrals from triarea: a cell is in the list of group ng iff its area survived the eps16
 * threshold (:2940-2948), in the reference's order (the order of the icellsd list: j outer, i inner) */
static void tri_lists(triangles *t, const blk *b)
{
    const size_t pe = (size_t)b->nxb * b->nyb;
    for (int ng = 1; ng <= NGROUPS; ++ng) {
        t->icells[ng - 1] = 0;
        for (int j = 1; j <= b->nyb; ++j)
            for (int i = 1; i <= b->nxb; ++i)
                if (G3(t->triarea, i, j, ng) != c0) {
                    const size_t ic = (size_t)(ng - 1) * pe + t->icells[ng - 1]++;
                    t->indxi[ic] = i; t->indxj[ic] = j;
                }
    }
}

int remap_oracle_triangles(int nxb, int nyb, int north, int coords, const double *dpx, const double *dpy,
                           const double *dxu, const double *dyu, double *xp, double *yp, int *iflux, int *jflux,
                           double *triarea, int *icells, double *edgearea)
{
    blk bb = {nxb, nyb, 2, nxb - 1, 2, nyb - 1};
    triangles t;
    tri_alloc(&t, &bb, xp, yp, triarea, iflux, jflux);
    if (!t.indxi || !t.indxj) return 2;
    const int rc = locate_triangles(&bb, north, dpx, dpy, dxu, dyu, &t, edgearea);
    if (rc == 0 && coords) triangle_coordinates(&bb, &t);
    for (int ng = 0; ng < NGROUPS; ++ng) icells[ng] = t.icells[ng];
    free(t.indxi); free(t.indxj);
    return rc;
}

int remap_oracle_transport_integrals(int nxb, int nyb, int ntrace, const int *tracer_type, const int *depend,
                                     const double *xp, const double *yp, const int *iflux, const int *jflux,
                                     const double *triarea, const double *mc, const double *mx, const double *my,
                                     double *mflx, const double *tc, const double *tx, const double *ty, double *mtflx)
{
    blk bb = {nxb, nyb, 2, nxb - 1, 2, nyb - 1};
    triangles t;
    tri_alloc(&t, &bb, (double *)xp, (double *)yp, (double *)triarea, (int *)iflux, (int *)jflux);
    if (!t.indxi || !t.indxj) return 2;
    tri_lists(&t, &bb);
    const int rc = transport_integrals(&bb, ntrace, tracer_type, depend, &t, mc, mx, my, mflx, tc, tx, ty, mtflx);
    free(t.indxi); free(t.indxj);
    return rc;
}

int remap_oracle_update_fields(int nxb, int nyb, int ntrace, const int *tracer_type, const int *depend,
                               const double *tarear, const double *mflxe, const double *mflxn, double *mm,
                               const double *mtflxe, const double *mtflxn, double *tm, int *l_stop, int *istop, int *jstop)
{
    blk bb = {nxb, nyb, 2, nxb - 1, 2, nyb - 1};
    *l_stop = 0; *istop = 0; *jstop = 0;
    return update_fields(&bb, ntrace, tracer_type, depend, tarear, l_stop, istop, jstop, mflxe, mflxn, mm, mtflxe, mtflxn, tm);
}

/* horizontal_remap (:309-850) on one global-with-ghost block: reconstruction, departure points, their ghost updates
 * (unless no_halo), then per edge (east, north) triangles + integration points + flux integrals of every category, and
 * update_fields of open water and of every ice category.  mm, tm are updated in place on the physical cells.
 * stop: 0 = completed; 1 = bad departure points (:552-563); 2 = negative open-water area; 3 = negative ice area, with
 * *stop_cat the category -- the reference aborts there; istop, jstop as it would print them. */
int remap_oracle_horizontal_remap(const evp_oracle_cfg *cfg, double dt, int ncat, int ntrace, const int *tracer_type,
                                  const int *depend, const int *has_dependents, int no_halo, const double *hm,
                                  const double *xav, const double *yav, const double *xxav, const double *yyav,
                                  const double *dxu, const double *dyu, const double *HTN, const double *HTE,
                                  const double *tarear, const double *uvel, const double *vvel, double *mm, double *tm,
                                  int *stop, int *stop_cat, int *istop, int *jstop)
{
    blk bb = {cfg->nx_global + 2, cfg->ny_global + 2, 2, cfg->nx_global + 1, 2, cfg->ny_global + 1};
    const blk *b = &bb;
    const size_t pe = (size_t)b->nxb * b->nyb, nct = (size_t)ncat * (ntrace ? ntrace : 1);
    *stop = 0; *stop_cat = 0; *istop = 0; *jstop = 0;
    double *mc = malloc(3 * (size_t)(ncat + 1) * pe * sizeof(double)), *tc = malloc(3 * nct * pe * sizeof(double));
    double *dp = malloc(2 * pe * sizeof(double));
    double *mflx = malloc(2 * (size_t)(ncat + 1) * pe * sizeof(double)), *mtflx = malloc(2 * nct * pe * sizeof(double));
    double *xp = malloc(2 * (size_t)NGROUPS * (NVERT + 1) * pe * sizeof(double)), *tri = malloc((size_t)NGROUPS * pe * sizeof(double));
    int *ifl = malloc(2 * (size_t)NGROUPS * pe * sizeof(int));
    if (!mc || !tc || !dp || !mflx || !mtflx || !xp || !tri || !ifl) return 2;
    double *mx = mc + (size_t)(ncat + 1) * pe, *my = mx + (size_t)(ncat + 1) * pe;
    double *tx = tc + nct * pe, *ty = tx + nct * pe, *dpx = dp, *dpy = dp + pe;
    double *yp = xp + (size_t)NGROUPS * (NVERT + 1) * pe;
    int *jfl = ifl + (size_t)NGROUPS * pe, rc, l_stop = 0;

    rc = remap_oracle_construct_fields(cfg, ncat, ntrace, tracer_type, depend, has_dependents, 1, hm, xav, yav, xxav, yyav,
                                       mm, tm, mc, mx, my, tc, tx, ty, NULL, NULL, NULL);
    if (rc) return rc;
    rc = remap_oracle_departure_points(cfg, dt, 1, 1, uvel, vvel, dxu, dyu, HTN, HTE, dpx, dpy, &l_stop, istop, jstop);
    if (rc) return rc;
    if (l_stop) { *stop = 1; goto done; }
    if (!no_halo) {                                                        /* :564-623 */
        if (evp_oracle_halo_r8(cfg, dpx, 1, 1, 0.0) || evp_oracle_halo_r8(cfg, dpy, 1, 1, 0.0)) return 1;
        for (int n = 0; n <= ncat; ++n)
            if (evp_oracle_halo_r8(cfg, mc + (size_t)n * pe, 0, 0, 0.0) || evp_oracle_halo_r8(cfg, mx + (size_t)n * pe, 0, 1, 0.0) ||
                evp_oracle_halo_r8(cfg, my + (size_t)n * pe, 0, 1, 0.0)) return 1;
        for (size_t k = 0; k < (size_t)ncat * ntrace; ++k)
            if (evp_oracle_halo_r8(cfg, tc + k * pe, 0, 0, 0.0) || evp_oracle_halo_r8(cfg, tx + k * pe, 0, 1, 0.0) ||
                evp_oracle_halo_r8(cfg, ty + k * pe, 0, 1, 0.0)) return 1;
    }
    for (int edge = 0; edge < 2; ++edge) {                                 /* east, then north (:668-770) */
        triangles t;
        tri_alloc(&t, b, xp, yp, tri, ifl, jfl);
        if (!t.indxi || !t.indxj) return 2;
        rc = locate_triangles(b, edge == 1, dpx, dpy, dxu, dyu, &t, NULL);
        if (rc) return rc;
        triangle_coordinates(b, &t);
        double *mf = mflx + (size_t)edge * (ncat + 1) * pe, *mtf = mtflx + (size_t)edge * nct * pe;
        rc = transport_integrals(b, ntrace, tracer_type, depend, &t, mc, mx, my, mf, NULL, NULL, NULL, NULL);
        for (int n = 1; n <= ncat && !rc; ++n) {
            const size_t o = (size_t)(n - 1) * ntrace * pe;
            rc = transport_integrals(b, ntrace, tracer_type, depend, &t, mc + (size_t)n * pe, mx + (size_t)n * pe,
                                     my + (size_t)n * pe, mf + (size_t)n * pe, ntrace ? tc + o : NULL, tx + o, ty + o, mtf + o);
        }
        free(t.indxi); free(t.indxj);
        if (rc) return rc;
    }
    {
        const double *mfe = mflx, *mfn = mflx + (size_t)(ncat + 1) * pe, *mtfe = mtflx, *mtfn = mtflx + nct * pe;
        rc = update_fields(b, ntrace, tracer_type, depend, tarear, &l_stop, istop, jstop, mfe, mfn, mm, NULL, NULL, NULL);
        if (rc) return rc;
        if (l_stop) { *stop = 2; goto done; }
        for (int n = 1; n <= ncat; ++n) {
            const size_t o = (size_t)(n - 1) * ntrace * pe;
            rc = update_fields(b, ntrace, tracer_type, depend, tarear, &l_stop, istop, jstop, mfe + (size_t)n * pe,
                               mfn + (size_t)n * pe, mm + (size_t)n * pe, mtfe + o, mtfn + o, ntrace ? tm + o : NULL);
            if (rc) return rc;
            if (l_stop) { *stop = 3; *stop_cat = n; goto done; }
        }
    }
done:
    free(mc); free(tc); free(dp); free(mflx); free(mtflx); free(xp); free(tri); free(ifl);
    return 0;
}

/* ---- the caller: transport_remap (source/ice_transport_driver.F90:198-627) with its compile-time settings
 * l_conservation_check = l_monotonicity_check = .false. (:249-251): state_to_tracers (:789-900), horizontal_remap,
 * tracers_to_state (:908-1002), bound_state (source/ice_state.F90:173-238, maskhalo_bound = .false.) ------------------ */
/* arrays: aicen, vicen, vsnon [ncat][nyb][nxb]; trcrn [ncat][ntrcr][nyb][nxb]; aim [ncat+1]...; trm [ncat][ntrace]... */
static void state_to_tracers(const blk *b, int ncat, int ntrcr, int nt_qsno, int nslyr, double rhos, double Lfresh,
                             const double *aice0, const double *aicen, const double *trcrn, const double *vicen,
                             const double *vsnon, double *aim, double *trm)
{
    const size_t pe = (size_t)b->nxb * b->nyb;
    const int ntrace = 2 + ntrcr;
    int *xi = malloc(pe * sizeof(int)), *xj = malloc(pe * sizeof(int));
    for (size_t q = 0; q < pe; ++q) aim[q] = aice0[q];
    for (int n = 1; n <= ncat; ++n) {
        double *am = aim + (size_t)n * pe, *tr = trm + (size_t)(n - 1) * ntrace * pe;
        const double *an = aicen + (size_t)(n - 1) * pe, *vi = vicen + (size_t)(n - 1) * pe, *vs = vsnon + (size_t)(n - 1) * pe;
        const double *tn = trcrn + (size_t)(n - 1) * ntrcr * pe;
        int icells = 0;
        for (size_t q = 0; q < (size_t)ntrace * pe; ++q) tr[q] = c0;
        for (int j = 1; j <= b->nyb; ++j)
            for (int i = 1; i <= b->nxb; ++i) {
                B2(am, i, j) = B2(an, i, j);
                if (B2(am, i, j) > puny) { xi[icells] = i; xj[icells] = j; ++icells; }
            }
        for (int ij = 0; ij < icells; ++ij) {
            const int i = xi[ij], j = xj[ij];
            const double w1 = c1 / B2(am, i, j);
            T3(tr, i, j, 1) = B2(vi, i, j) * w1;                           /* hice */
            T3(tr, i, j, 2) = B2(vs, i, j) * w1;                           /* hsno */
        }
        for (int it = 1; it <= ntrcr; ++it) {
            const int snow = it >= nt_qsno && it < nt_qsno + nslyr;        /* snow enthalpy, shifted to be positive */
            for (int ij = 0; ij < icells; ++ij) {
                const int i = xi[ij], j = xj[ij];
                T3(tr, i, j, 2 + it) = snow ? T3(tn, i, j, it) + rhos * Lfresh : T3(tn, i, j, it);
            }
        }
    }
    free(xi); free(xj);
}

static void tracers_to_state(const blk *b, int ncat, int ntrcr, int nt_qsno, int nslyr, double rhos, double Lfresh,
                             const double *aim, const double *trm, double *aice0, double *aicen, double *trcrn,
                             double *vicen, double *vsnon)
{
    const size_t pe = (size_t)b->nxb * b->nyb;
    const int ntrace = 2 + ntrcr;
    for (size_t q = 0; q < pe; ++q) aice0[q] = aim[q];
    for (int n = 1; n <= ncat; ++n) {
        const double *am = aim + (size_t)n * pe, *tr = trm + (size_t)(n - 1) * ntrace * pe;
        double *an = aicen + (size_t)(n - 1) * pe, *vi = vicen + (size_t)(n - 1) * pe, *vs = vsnon + (size_t)(n - 1) * pe;
        double *tn = trcrn + (size_t)(n - 1) * ntrcr * pe;
        for (int j = 1; j <= b->nyb; ++j)                                  /* cells with aim > 0 only: the others keep */
            for (int i = 1; i <= b->nxb; ++i) {                            /* their old state (:963-975)               */
                if (!(B2(am, i, j) > c0)) continue;
                B2(an, i, j) = B2(am, i, j);
                B2(vi, i, j) = B2(am, i, j) * T3(tr, i, j, 1);
                B2(vs, i, j) = B2(am, i, j) * T3(tr, i, j, 2);
                for (int it = 1; it <= ntrcr; ++it) {
                    const int snow = it >= nt_qsno && it < nt_qsno + nslyr;
                    T3(tn, i, j, it) = snow ? T3(tr, i, j, 2 + it) - rhos * Lfresh : T3(tr, i, j, 2 + it);
                }
            }
    }
}

int remap_oracle_state_to_tracers(int nxb, int nyb, int ncat, int ntrcr, int nt_qsno, int nslyr, double rhos, double Lfresh,
                                  const double *aice0, const double *aicen, const double *trcrn, const double *vicen,
                                  const double *vsnon, double *aim, double *trm)
{
    blk bb = {nxb, nyb, 2, nxb - 1, 2, nyb - 1};
    state_to_tracers(&bb, ncat, ntrcr, nt_qsno, nslyr, rhos, Lfresh, aice0, aicen, trcrn, vicen, vsnon, aim, trm);
    return 0;
}

int remap_oracle_tracers_to_state(int nxb, int nyb, int ncat, int ntrcr, int nt_qsno, int nslyr, double rhos, double Lfresh,
                                  const double *aim, const double *trm, double *aice0, double *aicen, double *trcrn,
                                  double *vicen, double *vsnon)
{
    blk bb = {nxb, nyb, 2, nxb - 1, 2, nyb - 1};
    tracers_to_state(&bb, ncat, ntrcr, nt_qsno, nslyr, rhos, Lfresh, aim, trm, aice0, aicen, trcrn, vicen, vsnon);
    return 0;
}

/* transport_remap on one global-with-ghost block; the state arrays are updated in place, ghost cells by bound_state
 * (unless no_halo).  stop etc. as remap_oracle_horizontal_remap (the reference aborts; the state is then left as read) */
int remap_oracle_transport_remap(const evp_oracle_cfg *cfg, double dt, int ncat, int ntrcr, int nt_qsno, int nslyr,
                                 double rhos, double Lfresh, const int *tracer_type, const int *depend,
                                 const int *has_dependents, int no_halo, const double *hm, const double *xav,
                                 const double *yav, const double *xxav, const double *yyav, const double *dxu,
                                 const double *dyu, const double *HTN, const double *HTE, const double *tarear,
                                 const double *uvel, const double *vvel, double *aice0, double *aicen, double *trcrn,
                                 double *vicen, double *vsnon, int *stop, int *stop_cat, int *istop, int *jstop)
{
    blk bb = {cfg->nx_global + 2, cfg->ny_global + 2, 2, cfg->nx_global + 1, 2, cfg->ny_global + 1};
    const size_t pe = (size_t)bb.nxb * bb.nyb;
    const int ntrace = 2 + ntrcr;
    double *aim = malloc((size_t)(ncat + 1) * pe * sizeof(double)), *trm = malloc((size_t)ncat * ntrace * pe * sizeof(double));
    if (!aim || !trm) return 2;
    state_to_tracers(&bb, ncat, ntrcr, nt_qsno, nslyr, rhos, Lfresh, aice0, aicen, trcrn, vicen, vsnon, aim, trm);
    int rc = remap_oracle_horizontal_remap(cfg, dt, ncat, ntrace, tracer_type, depend, has_dependents, no_halo, hm, xav, yav,
                                           xxav, yyav, dxu, dyu, HTN, HTE, tarear, uvel, vvel, aim, trm, stop, stop_cat, istop,
                                           jstop);
    if (rc == 0 && *stop == 0) {
        tracers_to_state(&bb, ncat, ntrcr, nt_qsno, nslyr, rhos, Lfresh, aim, trm, aice0, aicen, trcrn, vicen, vsnon);
        if (!no_halo) {                                                    /* bound_state */
            for (int n = 0; n < ncat && !rc; ++n) rc = evp_oracle_halo_r8(cfg, aicen + (size_t)n * pe, 0, 0, 0.0);
            for (int k = 0; k < ncat * ntrcr && !rc; ++k) rc = evp_oracle_halo_r8(cfg, trcrn + (size_t)k * pe, 0, 0, 0.0);
            for (int n = 0; n < ncat && !rc; ++n) rc = evp_oracle_halo_r8(cfg, vicen + (size_t)n * pe, 0, 0, 0.0);
            for (int n = 0; n < ncat && !rc; ++n) rc = evp_oracle_halo_r8(cfg, vsnon + (size_t)n * pe, 0, 0, 0.0);
        }
    }
    free(aim); free(trm);
    return rc;
}
