/*
 * evp_oracle.c -- CPU restatement of CICE5 evp(dt).  TEST INFRASTRUCTURE ONLY; see
 * evp_oracle.h for scope, citations and the parity note.
 *
 * Build: gcc -O2 -fopenmp -ffp-contract=off -fno-fast-math -shared -fPIC  (oracle/Makefile)
 * Evaluation order follows the Fortran statement by statement; x**2 is x*x.
 */
#include "evp_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------ */
/* constants, drivers/cice/ice_constants.F90:104-149 (computed, not literals) */
static const double c0 = 0.0, c1 = 1.0, c2 = 2.0, c4 = 4.0, p5 = 0.5, p25 = 0.25;
static const double puny = 1.0e-11;
static const double eyc = 0.36; /* ice_dyn_shared.F90:51 */

typedef struct {
    double p166, p333, p111, p055, p027, p222;
    /* set_evp_parameters, ice_dyn_shared.F90:209-257 (classic EVP branch) */
    double dte, dtei, ecci, dte2T, revp, arlx1i, brlx, denom1;
} Consts;

/* xmin = min(global_minval(dxt, tmask), global_minval(dyt, tmask)), ice_dyn_shared.F90:220-223;
 * only the revised-EVP branch uses it */
static void set_evp_parameters(Consts *k, double dt, int ndte, int revised_evp, double xmin)
{
    k->p166 = c1 / 6.0;
    k->p333 = c1 / 3.0;
    k->p111 = c1 / 9.0;
    k->p055 = k->p111 * p5;
    k->p027 = k->p055 * p5;
    k->p222 = c2 / 9.0;
    k->dte = dt / (double)ndte;          /* :209 */
    k->dtei = c1 / k->dte;               /* :210 */
    k->ecci = p25;                       /* :214 */
    {
        double tdamp2 = c2 * eyc * dt;   /* :217 */
        k->dte2T = k->dte / tdamp2;      /* :218 */
    }
    if (revised_evp) {                   /* Bouillon et al. 2013, :225-233 */
        const double Se = 0.86;          /* :226 */
        const double xi = 5.5e-3;        /* :227 */
        const double gamma = p25 * 1.e11 * dt;            /* :228 */
        k->revp = c1;                    /* :231 */
        k->arlx1i = c2 * xi / Se;        /* :232 */
        k->brlx = c2 * Se * xi * gamma / (xmin * xmin);   /* :233  xmin**2 */
    } else {
        k->revp = c0;                    /* :240 */
        k->arlx1i = k->dte2T;            /* :241 */
        k->brlx = dt * k->dtei;          /* :242 */
    }
    k->denom1 = c1 / (c1 + k->arlx1i);   /* :257 */
}

/* ------------------------------------------------------------------ */
/* decomposition: nslabs x-slabs, each one reference block with nghost = 1 */
typedef struct {
    int nx, ny;         /* global physical */
    int nslabs, nxl;    /* slabs, physical columns per slab */
    int nxb, nyb;       /* block dims incl. ghosts */
    int ilo, ihi, jlo, jhi;
    int ew_cyclic, ns_tripole;
    int no_halo;        /* unit-comparison mode: every halo update is skipped (one block, its ghost
                           cells stay as given) -- see evp_oracle_cfg.no_halo */
} Dom;

#define IX(d, i, j) ((size_t)((j) - 1) * (size_t)(d)->nxb + (size_t)((i) - 1))
#define GX(d, ig, jg) ((size_t)((jg) - 1) * (size_t)(d)->nx + (size_t)((ig) - 1))

/* global column index of local column i of slab b (ice_blocks.F90:266-309);
 * 0 means "no such column" (closed / open boundary) */
static int i_glob(const Dom *d, int b, int i)
{
    int ig = b * d->nxl + i - 1;
    if (ig < 1) ig = d->ew_cyclic ? ig + d->nx : 0;
    else if (ig > d->nx) ig = d->ew_cyclic ? ig - d->nx : 0;
    return ig;
}

static double **alloc_slabs(const Dom *d)
{
    double **f = (double **)malloc(sizeof(double *) * (size_t)d->nslabs);
    for (int b = 0; b < d->nslabs; ++b)
        f[b] = (double *)calloc((size_t)d->nxb * (size_t)d->nyb, sizeof(double));
    return f;
}
static int **alloc_slabs_i(const Dom *d)
{
    int **f = (int **)malloc(sizeof(int *) * (size_t)d->nslabs);
    for (int b = 0; b < d->nslabs; ++b)
        f[b] = (int *)calloc((size_t)d->nxb * (size_t)d->nyb, sizeof(int));
    return f;
}
static void free_slabs(const Dom *d, void **f)
{
    if (!f) return;
    for (int b = 0; b < d->nslabs; ++b) free(f[b]);
    free(f);
}

/* split a global-with-ghost array into slabs by column slicing */
static double **split_r8(const Dom *d, const double *g)
{
    double **f = alloc_slabs(d);
    const size_t gw = (size_t)d->nx + 2;
    if (!g) return f;
    for (int b = 0; b < d->nslabs; ++b)
        for (int j = 1; j <= d->nyb; ++j)
            memcpy(&f[b][IX(d, 1, j)], &g[(size_t)(j - 1) * gw + (size_t)b * d->nxl],
                   sizeof(double) * (size_t)d->nxb);
    return f;
}
static int **split_i4(const Dom *d, const int *g)
{
    int **f = alloc_slabs_i(d);
    const size_t gw = (size_t)d->nx + 2;
    if (!g) return f;
    for (int b = 0; b < d->nslabs; ++b)
        for (int j = 1; j <= d->nyb; ++j)
            memcpy(&f[b][IX(d, 1, j)], &g[(size_t)(j - 1) * gw + (size_t)b * d->nxl],
                   sizeof(int) * (size_t)d->nxb);
    return f;
}
/* merge: physical columns from the owner, W ghost from slab 0, E ghost from the last slab */
static void merge_r8(const Dom *d, double **f, double *g)
{
    const size_t gw = (size_t)d->nx + 2;
    if (!g) return;
    for (int b = 0; b < d->nslabs; ++b)
        for (int j = 1; j <= d->nyb; ++j) {
            int i0 = (b == 0) ? 1 : 2, i1 = (b == d->nslabs - 1) ? d->nxb : d->nxb - 1;
            memcpy(&g[(size_t)(j - 1) * gw + (size_t)b * d->nxl + (size_t)(i0 - 1)],
                   &f[b][IX(d, i0, j)], sizeof(double) * (size_t)(i1 - i0 + 1));
        }
}
static void merge_i4(const Dom *d, int **f, int *g)
{
    const size_t gw = (size_t)d->nx + 2;
    if (!g) return;
    for (int b = 0; b < d->nslabs; ++b)
        for (int j = 1; j <= d->nyb; ++j) {
            int i0 = (b == 0) ? 1 : 2, i1 = (b == d->nslabs - 1) ? d->nxb : d->nxb - 1;
            memcpy(&g[(size_t)(j - 1) * gw + (size_t)b * d->nxl + (size_t)(i0 - 1)],
                   &f[b][IX(d, i0, j)], sizeof(int) * (size_t)(i1 - i0 + 1));
        }
}

/* ------------------------------------------------------------------ */
/* ice_HaloUpdate, nghost = 1.
 *   - whole ghost ring pre-filled with `fill` (mpi/ice_boundary.F90:2625-2632)
 *   - E-W neighbour / cyclic copies of the physical rows (serial/ice_boundary.F90:706-725);
 *     S ghosts and (non-tripole) N ghosts have no source block -> stay `fill`
 *   - tripole u-fold through a global two-row buffer (serial/ice_boundary.F90:735-892;
 *     copy-in addresses :3737-3750, copy-out addresses :3763-3780)
 * loc: 0 center (ioffset=joffset=0), 1 NEcorner (ioffset=joffset=1 + seam averaging).
 * kind: 0 scalar (isign=+1), 1 vector (isign=-1).                              */
enum { LOC_CENTER = 0, LOC_NECORNER = 1 };
enum { KIND_SCALAR = 0, KIND_VECTOR = 1 };

/* first_pair: first column pair (i, nx - i) of the degenerate top row that the NE-corner fold averages.
 * The 2-D routine starts at i = 1 (mpi/ice_boundary.F90:1562, serial :817).  The 3-D routine of the MPI
 * build -- the one the velocity pair fld2 goes through (ice_dyn_evp.F90:314, :397) -- starts at
 * i = nxGlobal/2 - lenUsedTripoleBuf + 1 (mpi/ice_boundary.F90:2817), and ice_HaloCreate leaves
 * lenUsedTripoleBuf = nxGlobal/2 - 1 for a halo that spans the whole top row (:1067-1071, a length
 * taken as a difference of indices), so that loop starts at i = 2: the pair (1, nx - 1) next to the
 * pole is mirrored with its sign but NOT averaged.  Found by executing the reference text
 * (tests/golden/make_halo_vectors.py); reproduced here because parity is with the reference.        */
enum { FIRST_PAIR_2D = 1, FIRST_PAIR_3D = 2 };

static void halo_r8_ex(const Dom *d, double **f, int loc, int kind, double fill, int first_pair)
{
    const int nx = d->nx, ny = d->ny;
    double *G;
    if (d->no_halo) return;
    G = (double *)malloc(sizeof(double) * (size_t)nx * (size_t)ny);

    /* the sends: every block's physical cells */
    for (int b = 0; b < d->nslabs; ++b)
        for (int j = d->jlo; j <= d->jhi; ++j)
            for (int i = d->ilo; i <= d->ihi; ++i)
                G[GX(d, b * d->nxl + i - 1, j - 1)] = f[b][IX(d, i, j)];

    for (int b = 0; b < d->nslabs; ++b) {
        double *a = f[b];
        /* ring pre-fill */
        for (int i = 1; i <= d->nxb; ++i) {
            a[IX(d, i, 1)] = fill;
            a[IX(d, i, d->nyb)] = fill;
        }
        for (int j = 1; j <= d->nyb; ++j) {
            a[IX(d, 1, j)] = fill;
            a[IX(d, d->nxb, j)] = fill;
        }
        /* east / west neighbours (or cyclic wrap) */
        {
            int igw = i_glob(d, b, 1), ige = i_glob(d, b, d->nxb);
            for (int j = d->jlo; j <= d->jhi; ++j) {
                if (igw) a[IX(d, 1, j)] = G[GX(d, igw, j - 1)];
                if (ige) a[IX(d, d->nxb, j)] = G[GX(d, ige, j - 1)];
            }
        }
    }

    if (d->ns_tripole) {
        const int isign = (kind == KIND_SCALAR) ? 1 : -1;
        int ioffset, joffset;
        /* bufTripole(1:nx, 1:2) <- top two physical rows (:3737-3750) */
        double *buf1 = (double *)malloc(sizeof(double) * (size_t)nx); /* row ny-1 */
        double *buf2 = (double *)malloc(sizeof(double) * (size_t)nx); /* row ny   */
        for (int ig = 1; ig <= nx; ++ig) {
            buf1[ig - 1] = G[GX(d, ig, ny - 1)];
            buf2[ig - 1] = G[GX(d, ig, ny)];
        }
        if (loc == LOC_CENTER) {
            ioffset = 0; joffset = 0;                 /* :804-807 */
        } else {
            ioffset = 1; joffset = 1;                 /* :809-812 */
            for (int i = first_pair; i <= nx / 2 - 1; ++i) {   /* :817-824 */
                int iDst = nx - i;
                double x1 = buf2[i - 1], x2 = buf2[iDst - 1];
                double xavg = 0.5 * (x1 + isign * x2);
                buf2[i - 1] = xavg;
                buf2[iDst - 1] = isign * xavg;
            }
        }
        /* copy out, all local columns incl. E-W ghosts (:861-890) */
        for (int b = 0; b < d->nslabs; ++b) {
            for (int jj = 1; jj <= 2; ++jj) {
                for (int i = 1; i <= d->nxb; ++i) {
                    int ig = i_glob(d, b, i);
                    int iSrc, jSrc, jDst;
                    if (ig == 0) continue;            /* only cyclic E-W is meaningful here */
                    iSrc = nx - ig + 1;               /* :3768 */
                    jSrc = 1 + 3 - jj;                /* nghost + 3 - j, :3769 */
                    jDst = d->jhi + jj - 1;           /* :3776 */
                    iSrc -= ioffset;                  /* :874-877 */
                    jSrc -= joffset;
                    if (iSrc == 0) iSrc = nx;
                    if (iSrc > nx) iSrc -= nx;
                    if (jSrc <= 2 && jSrc > 0) {      /* :885 */
                        double v = (jSrc == 1) ? buf1[iSrc - 1] : buf2[iSrc - 1];
                        f[b][IX(d, i, jDst)] = isign * v;
                    }
                }
            }
        }
        free(buf1);
        free(buf2);
    }
    free(G);
}

static void halo_r8(const Dom *d, double **f, int loc, int kind, double fill)
{
    halo_r8_ex(d, f, loc, kind, fill, FIRST_PAIR_2D);
}

/* ice_HaloUpdate3DR8 of one level of fld2 (the velocity pair) */
static void halo_r8_fld2(const Dom *d, double **f, int loc, int kind, double fill)
{
    halo_r8_ex(d, f, loc, kind, fill, FIRST_PAIR_3D);
}

static void halo_i4(const Dom *d, int **f, int loc, int kind, int fill)
{
    double **t;
    if (d->no_halo) return;
    t = alloc_slabs(d);
    size_t n = (size_t)d->nxb * (size_t)d->nyb;
    for (int b = 0; b < d->nslabs; ++b)
        for (size_t q = 0; q < n; ++q) t[b][q] = (double)f[b][q];
    halo_r8(d, t, loc, kind, (double)fill);
    for (int b = 0; b < d->nslabs; ++b)
        for (size_t q = 0; q < n; ++q) f[b][q] = (int)t[b][q];
    free_slabs(d, (void **)t);
}

/* ice_HaloUpdate_stress(array1, array2, center, scalar): serial/ice_boundary.F90:3269-3443.
 * No ordinary neighbour copies, no ring pre-fill (mpi/ice_boundary.F90:5697-5710);
 * ghosts without a source block (the closed south edge) are set to fill = 0 (:3351-3353);
 * the north ghost row of array1 receives the mirrored top physical row of array2.   */
static void halo_stress(const Dom *d, double **a1, double **a2)
{
    const int nx = d->nx, ny = d->ny;
    double *buf2;
    if (d->no_halo) return;
    for (int b = 0; b < d->nslabs; ++b) {
        /* srcBlock == 0 entries: south row (and E/W ghosts when not cyclic) */
        for (int i = 1; i <= d->nxb; ++i) a1[b][IX(d, i, 1)] = 0.0;
        if (!d->ew_cyclic) {
            if (b == 0)
                for (int j = 1; j <= d->nyb; ++j) a1[b][IX(d, 1, j)] = 0.0;
            if (b == d->nslabs - 1)
                for (int j = 1; j <= d->nyb; ++j) a1[b][IX(d, d->nxb, j)] = 0.0;
        }
    }
    if (!d->ns_tripole) return;
    buf2 = (double *)malloc(sizeof(double) * (size_t)nx);
    for (int b = 0; b < d->nslabs; ++b)
        for (int i = d->ilo; i <= d->ihi; ++i)
            buf2[b * d->nxl + i - 2] = a2[b][IX(d, i, d->jhi)];
    (void)ny;
    for (int b = 0; b < d->nslabs; ++b)
        for (int i = 1; i <= d->nxb; ++i) {
            int ig = i_glob(d, b, i), iSrc;
            if (ig == 0) continue;
            iSrc = nx - ig + 1;                       /* ioffset = 0 (:3381) */
            if (iSrc == 0) iSrc = nx;
            /* jj = 2: jSrc = 2 -> ghost row jhi+1 ; jj = 1: jSrc = 3 > nghost+1, skipped */
            a1[b][IX(d, i, d->jhi + 1)] = buf2[iSrc - 1];
        }
    free(buf2);
}

/* ------------------------------------------------------------------ */
/* evp_prep1, ice_dyn_shared.F90:270-365 (one block) */
static void evp_prep1(const Dom *d, const evp_oracle_cfg *cfg,
                      const double *aice, const double *vice, const double *vsno,
                      const int *tmask, const double *strairxT, const double *strairyT,
                      double *strairx, double *strairy, double *tmass, int *icetmask)
{
    const size_t n = (size_t)d->nxb * (size_t)d->nyb;
    unsigned char *tmphm = (unsigned char *)malloc(n);
    for (int j = 1; j <= d->nyb; ++j)
        for (int i = 1; i <= d->nxb; ++i) {
            size_t q = IX(d, i, j);
            if (tmask[q])
                tmass[q] = (cfg->rhoi * vice[q] + cfg->rhos * vsno[q]);   /* :323 */
            else
                tmass[q] = c0;
            tmphm[q] = tmask[q] && (aice[q] > cfg->a_min) && (tmass[q] > cfg->m_min); /* :331 */
            strairx[q] = strairxT[q];                                     /* :339-340 */
            strairy[q] = strairyT[q];
            icetmask[q] = 0;                                              /* :345 */
        }
    for (int j = d->jlo; j <= d->jhi; ++j)
        for (int i = d->ilo; i <= d->ihi; ++i) {
            if (tmphm[IX(d, i - 1, j + 1)] || tmphm[IX(d, i, j + 1)] || tmphm[IX(d, i + 1, j + 1)] ||
                tmphm[IX(d, i - 1, j)] || tmphm[IX(d, i, j)] || tmphm[IX(d, i + 1, j)] ||
                tmphm[IX(d, i - 1, j - 1)] || tmphm[IX(d, i, j - 1)] || tmphm[IX(d, i + 1, j - 1)])
                icetmask[IX(d, i, j)] = 1;                                /* :354-358 */
            if (!tmask[IX(d, i, j)]) icetmask[IX(d, i, j)] = 0;           /* :360 */
        }
    free(tmphm);
}

/* to_ugrid, ice_grid.F90:1834-1876 (one block) */
static void to_ugrid(const Dom *d, const double *work1, double *work2,
                     const double *tarea, const double *uarea)
{
    memset(work2, 0, sizeof(double) * (size_t)d->nxb * (size_t)d->nyb);   /* :1853 */
    for (int j = d->jlo; j <= d->jhi; ++j)
        for (int i = d->ilo; i <= d->ihi; ++i)
            work2[IX(d, i, j)] = p25 *
                (work1[IX(d, i, j)] * tarea[IX(d, i, j)]
               + work1[IX(d, i + 1, j)] * tarea[IX(d, i + 1, j)]
               + work1[IX(d, i, j + 1)] * tarea[IX(d, i, j + 1)]
               + work1[IX(d, i + 1, j + 1)] * tarea[IX(d, i + 1, j + 1)])
               / uarea[IX(d, i, j)];                                      /* :1865-1870 */
}

/* to_tgrid, ice_grid.F90:1921-1958 (one block); work2 untouched outside physical cells */
static void to_tgrid(const Dom *d, const double *work1, double *work2,
                     const double *tarea, const double *uarea)
{
    for (int j = d->jlo; j <= d->jhi; ++j)
        for (int i = d->ilo; i <= d->ihi; ++i)
            work2[IX(d, i, j)] = p25 *
                (work1[IX(d, i, j)] * uarea[IX(d, i, j)]
               + work1[IX(d, i - 1, j)] * uarea[IX(d, i - 1, j)]
               + work1[IX(d, i, j - 1)] * uarea[IX(d, i, j - 1)]
               + work1[IX(d, i - 1, j - 1)] * uarea[IX(d, i - 1, j - 1)])
               / tarea[IX(d, i, j)];                                      /* :1947-1952 */
}

/* Fortran sign(c1, x): |1| with the sign of x, honouring -0 */
static inline double sign1(double x) { return copysign(1.0, x); }

/* evp_prep2, ice_dyn_shared.F90:377-614 (one block, classic EVP) */
typedef struct {
    int icellt, icellu;
    int *indxti, *indxtj, *indxui, *indxuj;
} Lists;

static void evp_prep2(const Dom *d, const evp_oracle_cfg *cfg, const Consts *k, Lists *L,
                      const double *aiu, const double *umass, double *umassdti,
                      const double *fcor, const int *umask, const double *uocn,
                      const double *vocn, const double *strairx, const double *strairy,
                      const double *ss_tltx, const double *ss_tlty, const int *icetmask,
                      int *iceumask, double *fm, double dt, double *strtltx, double *strtlty,
                      double *strocnx, double *strocny, double *strintx, double *strinty,
                      double *waterx, double *watery, double *forcex, double *forcey,
                      double *sp[4], double *sm[4], double *s12[4],
                      double *uvel_init, double *vvel_init, double *uvel, double *vvel)
{
    for (int j = 1; j <= d->nyb; ++j)
        for (int i = 1; i <= d->nxb; ++i) {
            size_t q = IX(d, i, j);
            waterx[q] = c0; watery[q] = c0; forcex[q] = c0; forcey[q] = c0;
            umassdti[q] = c0;                                             /* :486-490 */
            if (k->revp == 1.0 || icetmask[q] == 0)                       /* :492-518 */
                for (int c = 0; c < 4; ++c) { sp[c][q] = c0; sm[c][q] = c0; s12[c][q] = c0; }
        }
    L->icellt = 0;                                                        /* :528-537 */
    for (int j = d->jlo; j <= d->jhi + 1; ++j)
        for (int i = d->ilo; i <= d->ihi + 1; ++i)
            if (icetmask[IX(d, i, j)] == 1) {
                L->indxti[L->icellt] = i;
                L->indxtj[L->icellt] = j;
                L->icellt++;
            }
    L->icellu = 0;                                                        /* :545-577 */
    for (int j = d->jlo; j <= d->jhi; ++j)
        for (int i = d->ilo; i <= d->ihi; ++i) {
            size_t q = IX(d, i, j);
            int old = iceumask[q];
            iceumask[q] = umask[q] && (aiu[q] > cfg->a_min) && (umass[q] > cfg->m_min);
            if (iceumask[q]) {
                L->indxui[L->icellu] = i;
                L->indxuj[L->icellu] = j;
                L->icellu++;
                if (!old) { uvel[q] = uocn[q]; vvel[q] = vocn[q]; }       /* :560-563 */
            } else {
                uvel[q] = c0; vvel[q] = c0;                               /* :566-571 */
                strintx[q] = c0; strinty[q] = c0;
                strocnx[q] = c0; strocny[q] = c0;
            }
            uvel_init[q] = uvel[q];                                       /* :574-575 */
            vvel_init[q] = vvel[q];
        }
    for (int ij = 0; ij < L->icellu; ++ij) {                              /* :583-612 */
        size_t q = IX(d, L->indxui[ij], L->indxuj[ij]);
        umassdti[q] = umass[q] / dt;
        fm[q] = fcor[q] * umass[q];
        waterx[q] = uocn[q] * cfg->cosw - vocn[q] * cfg->sinw * sign1(fm[q]);
        watery[q] = vocn[q] * cfg->cosw + uocn[q] * cfg->sinw * sign1(fm[q]);
        if (cfg->tilt_mode == 0) {
            strtltx[q] = -fm[q] * vocn[q];                                /* :598-599 / :606-607 */
            strtlty[q] = fm[q] * uocn[q];
        } else {
            strtltx[q] = -cfg->gravit * umass[q] * ss_tltx[q];            /* :601-602 */
            strtlty[q] = -cfg->gravit * umass[q] * ss_tlty[q];
        }
        forcex[q] = strairx[q] + strtltx[q];
        forcey[q] = strairy[q] + strtlty[q];
    }
}

/* ------------------------------------------------------------------ */
/* stress, ice_dyn_evp.F90:520-849 (one block).  str is 8 planes of nxb*nyb. */
static void stress(const Dom *d, const Consts *k, int ksub, int ndte, const Lists *L,
                   const double *uvel, const double *vvel,
                   const double *dxt, const double *dyt, const double *dxhy, const double *dyhx,
                   const double *cxp, const double *cyp, const double *cxm, const double *cym,
                   const double *tarear, const double *tinyarea, const double *strength,
                   double *sp[4], double *sm[4], double *s12[4],
                   double *shear, double *divu, double *prs_sig,
                   double *rdg_conv, double *rdg_shear, double *str)
{
    const size_t n = (size_t)d->nxb * (size_t)d->nyb;
    const double ecci = k->ecci, arlx1i = k->arlx1i, denom1 = k->denom1;
    const double p027 = k->p027, p055 = k->p055, p111 = k->p111, p166 = k->p166,
                 p222 = k->p222, p333 = k->p333;
    double *stressp_1 = sp[0], *stressp_2 = sp[1], *stressp_3 = sp[2], *stressp_4 = sp[3];
    double *stressm_1 = sm[0], *stressm_2 = sm[1], *stressm_3 = sm[2], *stressm_4 = sm[3];
    double *stress12_1 = s12[0], *stress12_2 = s12[1], *stress12_3 = s12[2], *stress12_4 = s12[3];

    {   /* :613, str(:,:,:) = c0.  The reference does this once per block and its blocks run in parallel
         * (MPI tasks / OpenMP over iblk); with one block here, the fill is shared between the threads. */
        const size_t total = 8 * n, piece = (size_t)1 << 18;
        const long npieces = (long)((total + piece - 1) / piece);
#pragma omp parallel for schedule(static)
        for (long c = 0; c < npieces; ++c) {
            const size_t lo = (size_t)c * piece, len = (lo + piece <= total) ? piece : total - lo;
            memset(str + lo, 0, sizeof(double) * len);
        }
    }

#pragma omp parallel for schedule(static)
    for (int ij = 0; ij < L->icellt; ++ij) {
        const int i = L->indxti[ij], j = L->indxtj[ij];
        const size_t q = IX(d, i, j);
        const double u_ij = uvel[IX(d, i, j)], u_wj = uvel[IX(d, i - 1, j)];
        const double u_is = uvel[IX(d, i, j - 1)], u_ws = uvel[IX(d, i - 1, j - 1)];
        const double v_ij = vvel[IX(d, i, j)], v_wj = vvel[IX(d, i - 1, j)];
        const double v_is = vvel[IX(d, i, j - 1)], v_ws = vvel[IX(d, i - 1, j - 1)];
        double divune, divunw, divuse, divusw, tensionne, tensionnw, tensionse, tensionsw;
        double shearne, shearnw, shearse, shearsw, Deltane, Deltanw, Deltase, Deltasw;
        double c0ne, c0nw, c0se, c0sw, c1ne, c1nw, c1se, c1sw;
        double ssigpn, ssigps, ssigpe, ssigpw, ssigmn, ssigms, ssigme, ssigmw;
        double ssig12n, ssig12s, ssig12e, ssig12w, ssigp1, ssigp2, ssigm1, ssigm2, ssig121, ssig122;
        double csigpne, csigpnw, csigpse, csigpsw, csigmne, csigmnw, csigmse, csigmsw;
        double csig12ne, csig12nw, csig12se, csig12sw, str12ew, str12we, str12ns, str12sn;
        double strp_tmp, strm_tmp, tmp;

        /* strain rates * area (:627-654) */
        divune = cyp[q] * u_ij - dyt[q] * u_wj + cxp[q] * v_ij - dxt[q] * v_is;
        divunw = cym[q] * u_wj + dyt[q] * u_ij + cxp[q] * v_wj - dxt[q] * v_ws;
        divusw = cym[q] * u_ws + dyt[q] * u_is + cxm[q] * v_ws + dxt[q] * v_wj;
        divuse = cyp[q] * u_is - dyt[q] * u_ws + cxm[q] * v_is + dxt[q] * v_ij;

        tensionne = -cym[q] * u_ij - dyt[q] * u_wj + cxm[q] * v_ij + dxt[q] * v_is;
        tensionnw = -cyp[q] * u_wj + dyt[q] * u_ij + cxm[q] * v_wj + dxt[q] * v_ws;
        tensionsw = -cyp[q] * u_ws + dyt[q] * u_is + cxp[q] * v_ws - dxt[q] * v_wj;
        tensionse = -cym[q] * u_is - dyt[q] * u_ws + cxp[q] * v_is - dxt[q] * v_ij;

        shearne = -cym[q] * v_ij - dyt[q] * v_wj - cxm[q] * u_ij - dxt[q] * u_is;
        shearnw = -cyp[q] * v_wj + dyt[q] * v_ij - cxm[q] * u_wj - dxt[q] * u_ws;
        shearsw = -cyp[q] * v_ws + dyt[q] * v_is - cxp[q] * u_ws + dxt[q] * u_wj;
        shearse = -cym[q] * v_is - dyt[q] * v_ws - cxp[q] * u_is + dxt[q] * u_ij;

        /* Delta (:657-660) */
        Deltane = sqrt(divune * divune + ecci * (tensionne * tensionne + shearne * shearne));
        Deltanw = sqrt(divunw * divunw + ecci * (tensionnw * tensionnw + shearnw * shearnw));
        Deltase = sqrt(divuse * divuse + ecci * (tensionse * tensionse + shearse * shearse));
        Deltasw = sqrt(divusw * divusw + ecci * (tensionsw * tensionsw + shearsw * shearsw));

        if (ksub == ndte) {                                               /* :665-677 */
            divu[q] = p25 * (divune + divunw + divuse + divusw) * tarear[q];
            tmp = p25 * (Deltane + Deltanw + Deltase + Deltasw) * tarear[q];
            rdg_conv[q] = -fmin(divu[q], c0);
            rdg_shear[q] = p5 * (tmp - fabs(divu[q]));
            {
                double ts = tensionne + tensionnw + tensionse + tensionsw;
                double ss = shearne + shearnw + shearse + shearsw;
                shear[q] = p25 * tarear[q] * sqrt(ts * ts + ss * ss);
            }
        }

        /* replacement pressure / Delta (:683-697) */
        c0ne = strength[q] / fmax(Deltane, tinyarea[q]);
        c0nw = strength[q] / fmax(Deltanw, tinyarea[q]);
        c0sw = strength[q] / fmax(Deltasw, tinyarea[q]);
        c0se = strength[q] / fmax(Deltase, tinyarea[q]);
        prs_sig[q] = c0ne * Deltane;

        c1ne = c0ne * arlx1i; c1nw = c0nw * arlx1i; c1sw = c0sw * arlx1i; c1se = c0se * arlx1i;
        c0ne = c1ne * ecci;   c0nw = c1nw * ecci;   c0sw = c1sw * ecci;   c0se = c1se * ecci;

        /* the stresses (:704-721) */
        stressp_1[q] = (stressp_1[q] + c1ne * (divune - Deltane)) * denom1;
        stressp_2[q] = (stressp_2[q] + c1nw * (divunw - Deltanw)) * denom1;
        stressp_3[q] = (stressp_3[q] + c1sw * (divusw - Deltasw)) * denom1;
        stressp_4[q] = (stressp_4[q] + c1se * (divuse - Deltase)) * denom1;

        stressm_1[q] = (stressm_1[q] + c0ne * tensionne) * denom1;
        stressm_2[q] = (stressm_2[q] + c0nw * tensionnw) * denom1;
        stressm_3[q] = (stressm_3[q] + c0sw * tensionsw) * denom1;
        stressm_4[q] = (stressm_4[q] + c0se * tensionse) * denom1;

        stress12_1[q] = (stress12_1[q] + c0ne * shearne * p5) * denom1;
        stress12_2[q] = (stress12_2[q] + c0nw * shearnw * p5) * denom1;
        stress12_3[q] = (stress12_3[q] + c0sw * shearsw * p5) * denom1;
        stress12_4[q] = (stress12_4[q] + c0se * shearse * p5) * denom1;

        /* combinations (:752-795) */
        ssigpn = stressp_1[q] + stressp_2[q];
        ssigps = stressp_3[q] + stressp_4[q];
        ssigpe = stressp_1[q] + stressp_4[q];
        ssigpw = stressp_2[q] + stressp_3[q];
        ssigp1 = (stressp_1[q] + stressp_3[q]) * p055;
        ssigp2 = (stressp_2[q] + stressp_4[q]) * p055;

        ssigmn = stressm_1[q] + stressm_2[q];
        ssigms = stressm_3[q] + stressm_4[q];
        ssigme = stressm_1[q] + stressm_4[q];
        ssigmw = stressm_2[q] + stressm_3[q];
        ssigm1 = (stressm_1[q] + stressm_3[q]) * p055;
        ssigm2 = (stressm_2[q] + stressm_4[q]) * p055;

        ssig12n = stress12_1[q] + stress12_2[q];
        ssig12s = stress12_3[q] + stress12_4[q];
        ssig12e = stress12_1[q] + stress12_4[q];
        ssig12w = stress12_2[q] + stress12_3[q];
        ssig121 = (stress12_1[q] + stress12_3[q]) * p111;
        ssig122 = (stress12_2[q] + stress12_4[q]) * p111;

        csigpne = p111 * stressp_1[q] + ssigp2 + p027 * stressp_3[q];
        csigpnw = p111 * stressp_2[q] + ssigp1 + p027 * stressp_4[q];
        csigpsw = p111 * stressp_3[q] + ssigp2 + p027 * stressp_1[q];
        csigpse = p111 * stressp_4[q] + ssigp1 + p027 * stressp_2[q];

        csigmne = p111 * stressm_1[q] + ssigm2 + p027 * stressm_3[q];
        csigmnw = p111 * stressm_2[q] + ssigm1 + p027 * stressm_4[q];
        csigmsw = p111 * stressm_3[q] + ssigm2 + p027 * stressm_1[q];
        csigmse = p111 * stressm_4[q] + ssigm1 + p027 * stressm_2[q];

        csig12ne = p222 * stress12_1[q] + ssig122 + p055 * stress12_3[q];
        csig12nw = p222 * stress12_2[q] + ssig121 + p055 * stress12_4[q];
        csig12sw = p222 * stress12_3[q] + ssig122 + p055 * stress12_1[q];
        csig12se = p222 * stress12_4[q] + ssig121 + p055 * stress12_2[q];

        str12ew = p5 * dxt[q] * (p333 * ssig12e + p166 * ssig12w);
        str12we = p5 * dxt[q] * (p333 * ssig12w + p166 * ssig12e);
        str12ns = p5 * dyt[q] * (p333 * ssig12n + p166 * ssig12s);
        str12sn = p5 * dyt[q] * (p333 * ssig12s + p166 * ssig12n);

        /* dF/dx (:800-820) */
        strp_tmp = p25 * dyt[q] * (p333 * ssigpn + p166 * ssigps);
        strm_tmp = p25 * dyt[q] * (p333 * ssigmn + p166 * ssigms);
        str[0 * n + q] = -strp_tmp - strm_tmp - str12ew
                       + dxhy[q] * (-csigpne + csigmne) + dyhx[q] * csig12ne;
        str[1 * n + q] = strp_tmp + strm_tmp - str12we
                       + dxhy[q] * (-csigpnw + csigmnw) + dyhx[q] * csig12nw;
        strp_tmp = p25 * dyt[q] * (p333 * ssigps + p166 * ssigpn);
        strm_tmp = p25 * dyt[q] * (p333 * ssigms + p166 * ssigmn);
        str[2 * n + q] = -strp_tmp - strm_tmp + str12ew
                       + dxhy[q] * (-csigpse + csigmse) + dyhx[q] * csig12se;
        str[3 * n + q] = strp_tmp + strm_tmp + str12we
                       + dxhy[q] * (-csigpsw + csigmsw) + dyhx[q] * csig12sw;

        /* dF/dy (:825-845) */
        strp_tmp = p25 * dxt[q] * (p333 * ssigpe + p166 * ssigpw);
        strm_tmp = p25 * dxt[q] * (p333 * ssigme + p166 * ssigmw);
        str[4 * n + q] = -strp_tmp + strm_tmp - str12ns
                       - dyhx[q] * (csigpne + csigmne) + dxhy[q] * csig12ne;
        str[5 * n + q] = strp_tmp - strm_tmp - str12sn
                       - dyhx[q] * (csigpse + csigmse) + dxhy[q] * csig12se;
        strp_tmp = p25 * dxt[q] * (p333 * ssigpw + p166 * ssigpe);
        strm_tmp = p25 * dxt[q] * (p333 * ssigmw + p166 * ssigme);
        str[6 * n + q] = -strp_tmp + strm_tmp + str12ns
                       - dyhx[q] * (csigpnw + csigmnw) + dxhy[q] * csig12nw;
        str[7 * n + q] = strp_tmp - strm_tmp + str12sn
                       - dyhx[q] * (csigpsw + csigmsw) + dxhy[q] * csig12sw;
    }
}

/* stepu, ice_dyn_shared.F90:623-748 (one block) */
static void stepu(const Dom *d, const evp_oracle_cfg *cfg, const Consts *k, const Lists *L,
                  const double *Cw, const double *aiu, const double *str,
                  const double *uocn, const double *vocn, const double *waterx,
                  const double *watery, const double *forcex, const double *forcey,
                  const double *umassdti, const double *fm, const double *uarear,
                  double *strocnx, double *strocny, double *strintx, double *strinty,
                  const double *uvel_init, const double *vvel_init, double *uvel, double *vvel)
{
    const size_t n = (size_t)d->nxb * (size_t)d->nyb;
    const double rhow = cfg->rhow, cosw = cfg->cosw, sinw = cfg->sinw;
    const double brlx = k->brlx, revp = k->revp;
    /* NOTE: the reference updates uvel in place inside this loop; each iteration reads only
     * its own (i,j) velocity, so the parallel loop is equivalent. */
#pragma omp parallel for schedule(static)
    for (int ij = 0; ij < L->icellu; ++ij) {
        const int i = L->indxui[ij], j = L->indxuj[ij];
        const size_t q = IX(d, i, j);
        double uold = uvel[q], vold = vvel[q];
        double vrel, taux, tauy, cca, ccb, ab2, cc1, cc2;

        vrel = aiu[q] * rhow * Cw[q] * sqrt((uocn[q] - uold) * (uocn[q] - uold) +
                                            (vocn[q] - vold) * (vocn[q] - vold)); /* :708 */
        taux = vrel * waterx[q];
        tauy = vrel * watery[q];
        cca = (brlx + revp) * umassdti[q] + vrel * cosw;                  /* :715 */
        ccb = fm[q] + sign1(fm[q]) * vrel * sinw;                         /* :720 */
        ab2 = cca * cca + ccb * ccb;

        strintx[q] = uarear[q] *
            (str[0 * n + IX(d, i, j)] + str[1 * n + IX(d, i + 1, j)] +
             str[2 * n + IX(d, i, j + 1)] + str[3 * n + IX(d, i + 1, j + 1)]); /* :725 */
        strinty[q] = uarear[q] *
            (str[4 * n + IX(d, i, j)] + str[5 * n + IX(d, i, j + 1)] +
             str[6 * n + IX(d, i + 1, j)] + str[7 * n + IX(d, i + 1, j + 1)]); /* :727 */

        cc1 = strintx[q] + forcex[q] + taux + umassdti[q] * (brlx * uold + revp * uvel_init[q]);
        cc2 = strinty[q] + forcey[q] + tauy + umassdti[q] * (brlx * vold + revp * vvel_init[q]);

        uvel[q] = (cca * cc1 + ccb * cc2) / ab2;                          /* :736-737 */
        vvel[q] = (cca * cc2 - ccb * cc1) / ab2;

        strocnx[q] = taux;                                                /* :743-744 */
        strocny[q] = tauy;
    }
}

/* evp_finish, ice_dyn_shared.F90:757-844 (one block) */
static void evp_finish(const Dom *d, const evp_oracle_cfg *cfg, const Lists *L,
                       const double *Cw, const double *uvel, const double *vvel,
                       const double *uocn, const double *vocn, const double *aiu,
                       const double *fm, double *strocnx, double *strocny,
                       double *strocnxT, double *strocnyT)
{
    const double rhow = cfg->rhow, cosw = cfg->cosw, sinw = cfg->sinw;
    memset(strocnxT, 0, sizeof(double) * (size_t)d->nxb * (size_t)d->nyb); /* :806-811 */
    memset(strocnyT, 0, sizeof(double) * (size_t)d->nxb * (size_t)d->nyb);
    for (int ij = 0; ij < L->icellu; ++ij) {
        const size_t q = IX(d, L->indxui[ij], L->indxuj[ij]);
        double vrel = rhow * Cw[q] * sqrt((uocn[q] - uvel[q]) * (uocn[q] - uvel[q]) +
                                          (vocn[q] - vvel[q]) * (vocn[q] - vvel[q])); /* :818 */
        vrel = vrel * aiu[q];                                             /* :827 */
        strocnx[q] = vrel * ((uocn[q] - uvel[q]) * cosw - (vocn[q] - vvel[q]) * sinw * sign1(fm[q]));
        strocny[q] = vrel * ((vocn[q] - vvel[q]) * cosw + (uocn[q] - uvel[q]) * sinw * sign1(fm[q]));
        strocnxT[q] = strocnx[q] / aiu[q];                                /* :840-841 */
        strocnyT[q] = strocny[q] / aiu[q];
    }
}

#include "eap_oracle.inc"

/* ------------------------------------------------------------------ */
static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

#define NB (d->nslabs)
#define FOR_BLOCKS for (int b = 0; b < d->nslabs; ++b)

/* evp(dt), ice_dyn_evp.F90:68-510; with E != NULL eap(dt), ice_dyn_eap.F90:66-486 (the same driver around
 * stress_eap + stepa, no ice_HaloUpdate_stress) */
static int run_impl(const evp_oracle_cfg *cfg, const evp_oracle_grid *g,
                    evp_oracle_state *s, double dt, evp_oracle_stats *stats, evp_oracle_eap *E)
{
    Dom dom, *d = &dom;
    Consts K;
    double t_start = now_s(), t_loop0, t_loop1;
    size_t n;

    if (!cfg || !g || !s) return 1;
    if (cfg->nslabs < 1 || cfg->nx_global % cfg->nslabs) return 2;
    if (cfg->ndte < 1) return 3;
    if (cfg->ns_tripole && !cfg->ew_cyclic) return 4;
    d->nx = cfg->nx_global; d->ny = cfg->ny_global;
    d->nslabs = cfg->nslabs; d->nxl = d->nx / d->nslabs;
    d->nxb = d->nxl + 2; d->nyb = d->ny + 2;
    d->ilo = 2; d->ihi = d->nxl + 1; d->jlo = 2; d->jhi = d->ny + 1;
    d->ew_cyclic = cfg->ew_cyclic; d->ns_tripole = cfg->ns_tripole;
    d->no_halo = cfg->no_halo;
    if (d->no_halo && d->nslabs != 1) return 6;
    n = (size_t)d->nxb * (size_t)d->nyb;
    if (d->nxl < 2 || d->ny < 2) return 5;
#ifdef _OPENMP
    if (cfg->nthreads > 0) omp_set_num_threads(cfg->nthreads);
#endif
    {
        /* global_minval over the physical T cells with tmask (ice_global_reductions), :220-223 */
        double xmin = 1.0e300;
        const int gx = d->nx + 2;
        for (int j = 1; j <= d->ny; ++j)
            for (int i = 1; i <= d->nx; ++i) {
                const size_t q = (size_t)j * gx + i;
                if (g->tmask[q]) {
                    if (g->dxt[q] < xmin) xmin = g->dxt[q];
                    if (g->dyt[q] < xmin) xmin = g->dyt[q];
                }
            }
        set_evp_parameters(&K, cfg->dt_init, cfg->ndte, cfg->revised_evp, xmin);
    }

    /* ---- scatter to blocks ---- */
    double **dxt = split_r8(d, g->dxt), **dyt = split_r8(d, g->dyt);
    double **dxhy = split_r8(d, g->dxhy), **dyhx = split_r8(d, g->dyhx);
    double **cxp = split_r8(d, g->cxp), **cyp = split_r8(d, g->cyp);
    double **cxm = split_r8(d, g->cxm), **cym = split_r8(d, g->cym);
    double **tarea = split_r8(d, g->tarea), **uarea = split_r8(d, g->uarea);
    double **tarear = split_r8(d, g->tarear), **uarear = split_r8(d, g->uarear);
    double **tinyarea = split_r8(d, g->tinyarea), **fcor = split_r8(d, g->fcor);
    int **tmask = split_i4(d, g->tmask), **umask = split_i4(d, g->umask);

    double **aice = split_r8(d, s->aice), **vice = split_r8(d, s->vice);
    double **vsno = split_r8(d, s->vsno), **aice_init = split_r8(d, s->aice_init);
    double **strairxT = split_r8(d, s->strairxT), **strairyT = split_r8(d, s->strairyT);
    double **uocn = split_r8(d, s->uocn), **vocn = split_r8(d, s->vocn);
    double **ss_tltx = split_r8(d, s->ss_tltx), **ss_tlty = split_r8(d, s->ss_tlty);
    double **Cw = split_r8(d, s->Cw), **strength_in = split_r8(d, s->strength_in);

    double **uvel = split_r8(d, s->uvel), **vvel = split_r8(d, s->vvel);
    double **sp[4], **sm[4], **s12[4];
    for (int c = 0; c < 4; ++c) {
        sp[c] = split_r8(d, s->stressp[c]);
        sm[c] = split_r8(d, s->stressm[c]);
        s12[c] = split_r8(d, s->stress12[c]);
    }
    int **iceumask = split_i4(d, s->iceumask);
    double **fm = split_r8(d, s->fm);
    /* eap: structure tensor (inout), its cell averages (inout), strain / stress / yield diagnostics (out) */
    double **a11c[4] = {0, 0, 0, 0}, **a12c[4] = {0, 0, 0, 0}, **ea[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (E) {
        double *src[11] = {E->a11, E->a12, E->e11, E->e12, E->e22, E->s11, E->s12, E->s22,
                           E->yieldstress11, E->yieldstress12, E->yieldstress22};
        if (cfg->revised_evp) return 7;
        for (int c = 0; c < 4; ++c) { a11c[c] = split_r8(d, E->a11_c[c]); a12c[c] = split_r8(d, E->a12_c[c]); }
        for (int f = 0; f < 11; ++f) ea[f] = split_r8(d, src[f]);
    }

    /* outputs that the caller pre-zeroed via init_history_dyn (ice_flux.F90:639-673) are
     * taken as given so that values outside the ice lists pass through unchanged */
    double **strairx = split_r8(d, s->strairx), **strairy = split_r8(d, s->strairy);
    double **strtltx = split_r8(d, s->strtltx), **strtlty = split_r8(d, s->strtlty);
    double **strocnx = split_r8(d, s->strocnx), **strocny = split_r8(d, s->strocny);
    double **strintx = split_r8(d, s->strintx), **strinty = split_r8(d, s->strinty);
    double **strocnxT = split_r8(d, s->strocnxT), **strocnyT = split_r8(d, s->strocnyT);
    double **divu = split_r8(d, s->divu), **shear = split_r8(d, s->shear);
    double **rdg_conv = split_r8(d, s->rdg_conv), **rdg_shear = split_r8(d, s->rdg_shear);
    double **prs_sig = split_r8(d, s->prs_sig);
    double **uvel_init = alloc_slabs(d), **vvel_init = alloc_slabs(d);
    double **strength = alloc_slabs(d);

    /* evp locals (:120-137) */
    double **tmass = alloc_slabs(d), **waterx = alloc_slabs(d), **watery = alloc_slabs(d);
    double **forcex = alloc_slabs(d), **forcey = alloc_slabs(d), **aiu = alloc_slabs(d);
    double **umass = alloc_slabs(d), **umassdti = alloc_slabs(d), **work1 = alloc_slabs(d);
    int **icetmask = alloc_slabs_i(d);
    Lists *L = (Lists *)calloc((size_t)NB, sizeof(Lists));
    double *str = (double *)malloc(sizeof(double) * 8 * n);
    FOR_BLOCKS {
        L[b].indxti = (int *)malloc(sizeof(int) * n);
        L[b].indxtj = (int *)malloc(sizeof(int) * n);
        L[b].indxui = (int *)malloc(sizeof(int) * n);
        L[b].indxuj = (int *)malloc(sizeof(int) * n);
    }

    /* ---- :171-203 ---- */
    FOR_BLOCKS {
        memset(rdg_conv[b], 0, sizeof(double) * n);
        memset(rdg_shear[b], 0, sizeof(double) * n);
        memset(divu[b], 0, sizeof(double) * n);
        memset(shear[b], 0, sizeof(double) * n);
        memset(prs_sig[b], 0, sizeof(double) * n);
        if (E) for (int f = 2; f < 11; ++f) memset(ea[f][b], 0, sizeof(double) * n);   /* ice_dyn_eap.F90:174-182 */
        evp_prep1(d, cfg, aice[b], vice[b], vsno[b], tmask[b], strairxT[b], strairyT[b],
                  strairx[b], strairy[b], tmass[b], icetmask[b]);
    }
    /* :210 */
    halo_i4(d, icetmask, LOC_CENTER, KIND_SCALAR, 0);
    /* :218-219 */
    FOR_BLOCKS {
        to_ugrid(d, tmass[b], umass[b], tarea[b], uarea[b]);
        to_ugrid(d, aice_init[b], aiu[b], tarea[b], uarea[b]);
    }
    if (s->strax && s->stray) {
        /* :220-229 (#ifdef ACCESS, .not. calc_strair): strairx(:,:,:) = strax(:,:,:), ghosts included */
        double **strax = split_r8(d, s->strax), **stray = split_r8(d, s->stray);
        FOR_BLOCKS memcpy(strairx[b], strax[b], sizeof(double) * n);
        FOR_BLOCKS memcpy(strairy[b], stray[b], sizeof(double) * n);
        free_slabs(d, (void **)strax);
        free_slabs(d, (void **)stray);
    } else {
        /* :240-241 t2ugrid_vector (ice_grid.F90:1799-1823) */
        FOR_BLOCKS memcpy(work1[b], strairx[b], sizeof(double) * n);
        halo_r8(d, work1, LOC_CENTER, KIND_VECTOR, 0.0);
        FOR_BLOCKS to_ugrid(d, work1[b], strairx[b], tarea[b], uarea[b]);
        FOR_BLOCKS memcpy(work1[b], strairy[b], sizeof(double) * n);
        halo_r8(d, work1, LOC_CENTER, KIND_VECTOR, 0.0);
        FOR_BLOCKS to_ugrid(d, work1[b], strairy[b], tarea[b], uarea[b]);
    }

    /* ---- :247-308 ---- */
    FOR_BLOCKS {
        double *spb[4], *smb[4], *s12b[4];
        for (int c = 0; c < 4; ++c) { spb[c] = sp[c][b]; smb[c] = sm[c][b]; s12b[c] = s12[c][b]; }
        evp_prep2(d, cfg, &K, &L[b], aiu[b], umass[b], umassdti[b], fcor[b], umask[b],
                  uocn[b], vocn[b], strairx[b], strairy[b], ss_tltx[b], ss_tlty[b],
                  icetmask[b], iceumask[b], fm[b], dt, strtltx[b], strtlty[b],
                  strocnx[b], strocny[b], strintx[b], strinty[b], waterx[b], watery[b],
                  forcex[b], forcey[b], spb, smb, s12b, uvel_init[b], vvel_init[b],
                  uvel[b], vvel[b]);
        if (E) {                                        /* ice_dyn_eap.F90:274-289: every cell of the block */
            for (size_t q = 0; q < n; ++q)
                if (icetmask[b][q] == 0)
                    for (int c = 0; c < 4; ++c) { a11c[c][b][q] = p5; a12c[c][b][q] = c0; }
        }
        /* ice_strength (:291): host-supplied values; 0 outside the block interior / list */
        memset(strength[b], 0, sizeof(double) * n);
        if (cfg->strength_masked) {
            for (int ij = 0; ij < L[b].icellt; ++ij) {
                size_t q = IX(d, L[b].indxti[ij], L[b].indxtj[ij]);
                strength[b][q] = strength_in[b][q];
            }
        } else {
            for (int j = d->jlo; j <= d->jhi; ++j)
                for (int i = d->ilo; i <= d->ihi; ++i)
                    strength[b][IX(d, i, j)] = strength_in[b][IX(d, i, j)];
        }
    }
    /* :311-315 */
    halo_r8(d, strength, LOC_CENTER, KIND_SCALAR, 0.0);
    halo_r8_fld2(d, uvel, LOC_NECORNER, KIND_VECTOR, 0.0);            /* fld2: the 3-D routine */
    halo_r8_fld2(d, vvel, LOC_NECORNER, KIND_VECTOR, 0.0);

    /* ---- subcycling :336-410 ---- */
    t_loop0 = now_s();
    for (int ksub = 1; ksub <= cfg->ndte; ++ksub) {
        FOR_BLOCKS {
            double *spb[4], *smb[4], *s12b[4];
            for (int c = 0; c < 4; ++c) { spb[c] = sp[c][b]; smb[c] = sm[c][b]; s12b[c] = s12[c][b]; }
            double *a11b[4] = {0, 0, 0, 0}, *a12b[4] = {0, 0, 0, 0};
            if (E) {
                for (int c = 0; c < 4; ++c) { a11b[c] = a11c[c][b]; a12b[c] = a12c[c][b]; }
                stress_eap(d, &K, ksub, cfg->ndte, &L[b], E, uvel[b], vvel[b], dxt[b], dyt[b], dxhy[b], dyhx[b],
                           cxp[b], cyp[b], cxm[b], cym[b], tarear[b], strength[b], a11b, a12b, spb, smb, s12b,
                           shear[b], divu[b], ea[2][b], ea[3][b], ea[4][b], ea[5][b], ea[6][b], ea[7][b],
                           ea[8][b], ea[9][b], ea[10][b], prs_sig[b], rdg_conv[b], str);
            } else
            stress(d, &K, ksub, cfg->ndte, &L[b], uvel[b], vvel[b], dxt[b], dyt[b], dxhy[b],
                   dyhx[b], cxp[b], cyp[b], cxm[b], cym[b], tarear[b], tinyarea[b],
                   strength[b], spb, smb, s12b, shear[b], divu[b], prs_sig[b],
                   rdg_conv[b], rdg_shear[b], str);
            stepu(d, cfg, &K, &L[b], Cw[b], aiu[b], str, uocn[b], vocn[b], waterx[b],
                  watery[b], forcex[b], forcey[b], umassdti[b], fm[b], uarear[b],
                  strocnx[b], strocny[b], strintx[b], strinty[b], uvel_init[b],
                  vvel_init[b], uvel[b], vvel[b]);
            if (E && ksub % 10 == 1)                    /* ice_dyn_eap.F90:408-421: only every 10th subcycle */
                stepa(d, E->trig_mode, K.dtei, &L[b], ea[0][b], ea[1][b], a11b, a12b, spb, smb, s12b);
        }
        halo_r8_fld2(d, uvel, LOC_NECORNER, KIND_VECTOR, 0.0);            /* :397 */
        halo_r8_fld2(d, vvel, LOC_NECORNER, KIND_VECTOR, 0.0);
    }
    t_loop1 = now_s();

    /* ---- tripole stress symmetrisation :416-481 ---- */
    if (cfg->grid_tripole && !E) {                      /* eap has no stress halo */
        halo_stress(d, sp[0], sp[2]);  halo_stress(d, sp[2], sp[0]);
        halo_stress(d, sp[1], sp[3]);  halo_stress(d, sp[3], sp[1]);
        halo_stress(d, sm[0], sm[2]);  halo_stress(d, sm[2], sm[0]);
        halo_stress(d, sm[1], sm[3]);  halo_stress(d, sm[3], sm[1]);
        halo_stress(d, s12[0], s12[2]); halo_stress(d, s12[2], s12[0]);
        halo_stress(d, s12[1], s12[3]); halo_stress(d, s12[3], s12[1]);
    }

    /* ---- :487-506 ---- */
    FOR_BLOCKS evp_finish(d, cfg, &L[b], Cw[b], uvel[b], vvel[b], uocn[b], vocn[b], aiu[b],
                          fm[b], strocnx[b], strocny[b], strocnxT[b], strocnyT[b]);
    FOR_BLOCKS memcpy(work1[b], strocnxT[b], sizeof(double) * n);
    halo_r8(d, work1, LOC_NECORNER, KIND_VECTOR, 0.0);
    FOR_BLOCKS to_tgrid(d, work1[b], strocnxT[b], tarea[b], uarea[b]);
    FOR_BLOCKS memcpy(work1[b], strocnyT[b], sizeof(double) * n);
    halo_r8(d, work1, LOC_NECORNER, KIND_VECTOR, 0.0);
    FOR_BLOCKS to_tgrid(d, work1[b], strocnyT[b], tarea[b], uarea[b]);

    if (stats) {
        stats->loop_seconds = t_loop1 - t_loop0;
        stats->active_t_cells = 0;
        stats->active_u_cells = 0;
        FOR_BLOCKS { stats->active_t_cells += L[b].icellt; stats->active_u_cells += L[b].icellu; }
    }

    /* ---- gather results ---- */
    merge_r8(d, uvel, s->uvel); merge_r8(d, vvel, s->vvel);
    for (int c = 0; c < 4; ++c) {
        merge_r8(d, sp[c], s->stressp[c]);
        merge_r8(d, sm[c], s->stressm[c]);
        merge_r8(d, s12[c], s->stress12[c]);
    }
    merge_i4(d, iceumask, s->iceumask); merge_r8(d, fm, s->fm);
    merge_r8(d, strairx, s->strairx); merge_r8(d, strairy, s->strairy);
    merge_r8(d, strtltx, s->strtltx); merge_r8(d, strtlty, s->strtlty);
    merge_r8(d, strocnx, s->strocnx); merge_r8(d, strocny, s->strocny);
    merge_r8(d, strintx, s->strintx); merge_r8(d, strinty, s->strinty);
    merge_r8(d, strocnxT, s->strocnxT); merge_r8(d, strocnyT, s->strocnyT);
    merge_r8(d, divu, s->divu); merge_r8(d, shear, s->shear);
    merge_r8(d, rdg_conv, s->rdg_conv); merge_r8(d, rdg_shear, s->rdg_shear);
    merge_r8(d, prs_sig, s->prs_sig);
    merge_r8(d, uvel_init, s->uvel_init); merge_r8(d, vvel_init, s->vvel_init);
    merge_r8(d, tmass, s->tmass); merge_r8(d, strength, s->strength);
    merge_i4(d, icetmask, s->icetmask);
    if (E) {
        double *dst[11] = {E->a11, E->a12, E->e11, E->e12, E->e22, E->s11, E->s12, E->s22,
                           E->yieldstress11, E->yieldstress12, E->yieldstress22};
        for (int c = 0; c < 4; ++c) {
            merge_r8(d, a11c[c], E->a11_c[c]); merge_r8(d, a12c[c], E->a12_c[c]);
            free_slabs(d, (void **)a11c[c]); free_slabs(d, (void **)a12c[c]);
        }
        for (int f = 0; f < 11; ++f) { merge_r8(d, ea[f], dst[f]); free_slabs(d, (void **)ea[f]); }
    }

    /* ---- free ---- */
    {
        double **all[] = { dxt, dyt, dxhy, dyhx, cxp, cyp, cxm, cym, tarea, uarea, tarear, uarear,
            tinyarea, fcor, aice, vice, vsno, aice_init, strairxT, strairyT, uocn, vocn, ss_tltx,
            ss_tlty, Cw, strength_in, uvel, vvel, fm, strairx, strairy, strtltx, strtlty, strocnx,
            strocny, strintx, strinty, strocnxT, strocnyT, divu, shear, rdg_conv, rdg_shear,
            prs_sig, uvel_init, vvel_init, strength, tmass, waterx, watery, forcex, forcey, aiu,
            umass, umassdti, work1, sp[0], sp[1], sp[2], sp[3], sm[0], sm[1], sm[2], sm[3],
            s12[0], s12[1], s12[2], s12[3] };
        for (size_t q = 0; q < sizeof(all) / sizeof(all[0]); ++q) free_slabs(d, (void **)all[q]);
        free_slabs(d, (void **)tmask); free_slabs(d, (void **)umask);
        free_slabs(d, (void **)iceumask); free_slabs(d, (void **)icetmask);
        FOR_BLOCKS { free(L[b].indxti); free(L[b].indxtj); free(L[b].indxui); free(L[b].indxuj); }
        free(L); free(str);
    }
    if (stats) stats->total_seconds = now_s() - t_start;
    return 0;
}

int evp_oracle_run(const evp_oracle_cfg *cfg, const evp_oracle_grid *g,
                   evp_oracle_state *s, double dt, evp_oracle_stats *stats)
{
    return run_impl(cfg, g, s, dt, stats, NULL);
}

int eap_oracle_run(const evp_oracle_cfg *cfg, const evp_oracle_grid *g, evp_oracle_state *s, evp_oracle_eap *e,
                   double dt, evp_oracle_stats *stats)
{
    if (!e) return 1;
    return run_impl(cfg, g, s, dt, stats, e);
}

/* ------------------------------------------------------------------ */
static void cfg_to_dom(const evp_oracle_cfg *cfg, Dom *d)
{
    d->nx = cfg->nx_global; d->ny = cfg->ny_global;
    d->nslabs = cfg->nslabs > 0 ? cfg->nslabs : 1; d->nxl = d->nx / d->nslabs;
    d->nxb = d->nxl + 2; d->nyb = d->ny + 2;
    d->ilo = 2; d->ihi = d->nxl + 1; d->jlo = 2; d->jhi = d->ny + 1;
    d->ew_cyclic = cfg->ew_cyclic; d->ns_tripole = cfg->ns_tripole;
    d->no_halo = 0;
}

int evp_oracle_halo_r8(const evp_oracle_cfg *cfg, double *a, int loc, int kind, double fill)
{
    Dom dom; double **f;
    if (!cfg || !a || cfg->nx_global % (cfg->nslabs > 0 ? cfg->nslabs : 1)) return 1;
    cfg_to_dom(cfg, &dom);
    f = split_r8(&dom, a);
    halo_r8(&dom, f, loc, kind, fill);
    merge_r8(&dom, f, a);
    free_slabs(&dom, (void **)f);
    return 0;
}

int evp_oracle_halo_r8_fld2(const evp_oracle_cfg *cfg, double *a, int loc, int kind, double fill)
{
    Dom dom; double **f;
    if (!cfg || !a || cfg->nx_global % (cfg->nslabs > 0 ? cfg->nslabs : 1)) return 1;
    cfg_to_dom(cfg, &dom);
    f = split_r8(&dom, a);
    halo_r8_fld2(&dom, f, loc, kind, fill);
    merge_r8(&dom, f, a);
    free_slabs(&dom, (void **)f);
    return 0;
}

int evp_oracle_halo_i4(const evp_oracle_cfg *cfg, int *a, int loc, int kind, int fill)
{
    Dom dom; int **f;
    if (!cfg || !a || cfg->nx_global % (cfg->nslabs > 0 ? cfg->nslabs : 1)) return 1;
    cfg_to_dom(cfg, &dom);
    f = split_i4(&dom, a);
    halo_i4(&dom, f, loc, kind, fill);
    merge_i4(&dom, f, a);
    free_slabs(&dom, (void **)f);
    return 0;
}

int evp_oracle_halo_stress(const evp_oracle_cfg *cfg, double *a1, const double *a2)
{
    Dom dom; double **f1, **f2;
    if (!cfg || !a1 || !a2 || cfg->nx_global % (cfg->nslabs > 0 ? cfg->nslabs : 1)) return 1;
    cfg_to_dom(cfg, &dom);
    f1 = split_r8(&dom, a1);
    f2 = split_r8(&dom, a2);
    halo_stress(&dom, f1, f2);
    merge_r8(&dom, f1, a1);
    free_slabs(&dom, (void **)f1);
    free_slabs(&dom, (void **)f2);
    return 0;
}

/* What restartfile does to the 12 stresses once they are read (source/ice_restart_driver.F90): on a tripole
 * grid the 12 ice_HaloUpdate_stress calls (:370-394, the pairing of ice_dyn_evp.F90:416-481), then the unused
 * south ghost row and west ghost column of every block are zeroed (:431-466, j = 1..nghost / i = 1..nghost).
 * s[12]: stressp_1..4, stressm_1..4, stress12_1..4 as global-with-ghost planes, in place. */
int evp_oracle_restart_fixup(const evp_oracle_cfg *cfg, double *const *s)
{
    Dom dom; double **f[12];
    if (!cfg || !s || cfg->nx_global % (cfg->nslabs > 0 ? cfg->nslabs : 1)) return 1;
    cfg_to_dom(cfg, &dom);
    for (int q = 0; q < 12; ++q) f[q] = split_r8(&dom, s[q]);
    if (cfg->grid_tripole) {
        for (int g = 0; g < 3; ++g) {
            double ***b = f + 4 * g;
            halo_stress(&dom, b[0], b[2]);  halo_stress(&dom, b[2], b[0]);
            halo_stress(&dom, b[1], b[3]);  halo_stress(&dom, b[3], b[1]);
        }
    }
    for (int q = 0; q < 12; ++q)
        for (int b = 0; b < dom.nslabs; ++b) {
            for (int i = 1; i <= dom.nxb; ++i) f[q][b][IX(&dom, i, 1)] = 0.0;
            for (int j = 1; j <= dom.nyb; ++j) f[q][b][IX(&dom, 1, j)] = 0.0;
        }
    for (int q = 0; q < 12; ++q) { merge_r8(&dom, f[q], s[q]); free_slabs(&dom, (void **)f[q]); }
    return 0;
}

/* principal_stress, ice_dyn_shared.F90:853-893; spval_dbl = 1e30 (ice_constants.F90) */
void evp_oracle_principal_stress(long long n, const double *stressp_1, const double *stressm_1,
                                 const double *stress12_1, const double *prs_sig,
                                 double *sig1, double *sig2)
{
    const double spval_dbl = 1.0e30;
    for (long long q = 0; q < n; ++q) {
        if (prs_sig[q] > puny) {
            double r = sqrt(stressm_1[q] * stressm_1[q] + c4 * (stress12_1[q] * stress12_1[q]));
            sig1[q] = (p5 * (stressp_1[q] + r)) / prs_sig[q];
            sig2[q] = (p5 * (stressp_1[q] - r)) / prs_sig[q];
        } else {
            sig1[q] = spval_dbl;
            sig2[q] = spval_dbl;
        }
    }
}
