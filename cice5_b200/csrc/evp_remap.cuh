// evp_remap.cuh -- first stages of horizontal transport on the device: departure_points and the
// field reconstruction (make_masks + construct_fields + limited_gradient, second half of this file).
//
// reference: source/ice_transport_remap.F90:1493-1670 (departure_points, nghost = 1), called by horizontal_remap
// (:541-550) right after evp in step_dynamics (source/ice_step_mod.F90:1126-1130); the ghost cells of dpx, dpy are then
// filled by two ice_HaloUpdate calls (NE corner, vector; :574-578) -- the generic halo of evp_prep.cuh.  Reads uvel,
// vvel straight from the resident EVP state: the velocities never leave the device between evp and this stage.
#pragma once
#include "evp_kernels.cuh"

namespace evp {

struct DepartArgs {
    int nxb, nyb, ld;              // block incl. ghosts, plane pitch
    double dt;
    const double *u, *v;           // resident velocities (current ping-pong copy)
    const double *dxu, *dyu, *htn, *hte;
    double *dpx, *dpy;
    int *stop;                     // 0: all departure points in bounds; else 1 + (j-1)*nxb + (i-1) of the LAST
                                   // offending cell in the reference's loop order (j outer, i inner; 1-based i, j)
    int midpt;                     // l_dp_midpt
};

// dpx = -dt*uvel, dpy = -dt*vvel on the physical cells, 0 on the ghost ring; bounds check (:1557-1575)
__global__ void departure_first_kernel(const __grid_constant__ DepartArgs a)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c >= a.nxb) return;
    const size_t q = (size_t)r * a.ld + c;
    double px = 0.0, py = 0.0;
    if (c >= 1 && c <= a.nxb - 2 && r >= 1 && r <= a.nyb - 2) {
        px = -a.dt * a.u[q];
        py = -a.dt * a.v[q];
        if (px < -a.htn[q] || px > a.htn[q + 1] || py < -a.hte[q] || py > a.hte[q + a.ld])
            atomicMax(a.stop, 1 + r * a.nxb + c);
    }
    a.dpx[q] = px;
    a.dpy[q] = py;
}

// midpoint correction (:1593-1668): skipped altogether when any cell was out of bounds (the reference returns
// before it) and for cells at rest
__global__ void departure_midpt_kernel(const __grid_constant__ DepartArgs a)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (!a.midpt || c < 1 || c > a.nxb - 2 || r < 1 || r > a.nyb - 2) return;
    if (*(volatile int *)a.stop) return;
    const size_t q = (size_t)r * a.ld + c;
    if (a.u[q] == 0.0 && a.v[q] == 0.0) return;
    const double sx = a.dpx[q] / a.dxu[q], sy = a.dpy[q] / a.dyu[q];        // unit-cell coordinates
    const double mpx = 0.5 * sx, mpy = 0.5 * sy;                             // midpoint of the back trajectory
    // cell (i2, j2) that holds the midpoint, and the midpoint relative to that cell's centre
    const int di = mpx >= 0.0 ? 1 : 0, dj = mpy >= 0.0 ? 1 : 0;
    const double mpxt = mpx >= 0.0 ? mpx - 0.5 : mpx + 0.5;
    const double mpyt = mpy >= 0.0 ? mpy - 0.5 : mpy + 0.5;
    const size_t q2 = q + (size_t)dj * a.ld + di;                            // (i2, j2)
    const double xm = mpxt - 0.5, xp = mpxt + 0.5, ym = mpyt - 0.5, yp = mpyt + 0.5;
    const double ump = a.u[q2 - a.ld - 1] * xm * ym - a.u[q2 - a.ld] * xp * ym + a.u[q2] * xp * yp - a.u[q2 - 1] * xm * yp;
    const double vmp = a.v[q2 - a.ld - 1] * xm * ym - a.v[q2 - a.ld] * xp * ym + a.v[q2] * xp * yp - a.v[q2 - 1] * xm * yp;
    a.dpx[q] = -a.dt * ump;
    a.dpy[q] = -a.dt * vmp;
}


// ---- second stage: make_masks + construct_fields + limited_gradient, one pass -----------------------------------
// reference: source/ice_transport_remap.F90:867-1015 (make_masks), :1024-1340 (construct_fields), :1348-1484
// (limited_gradient), called per block by horizontal_remap (:462-511).  The reference walks index lists and keeps
// mmask, tmask, mxav, myav, mtxav, mtyav in memory; every one of them is a function of values within one cell of the
// cell it is used in, so here one thread owns one cell of one category and carries them in registers: mm and the
// tracer planes are read once (3 x 3 neighbourhoods out of L1), mc mx my tc tx ty written once, nothing else touches
// memory.  Each expression is the reference's, operand for operand (the library is built without FMA contraction);
// min / max pick the first smallest / largest argument as a Fortran processor evaluating the list left to right.
constexpr int kRemapMaxTrace = 128;    // tracers per category (EVP_REMAP_MAX_TRACE)
constexpr int kRemapMaxParents = 16;   // type-1 tracers with dependents (hice, hsno, apnd, ...)

struct ReconArgs {
    int nxb, nyb, ld;
    size_t pe;                         // plane stride (elements)
    int ncat, ntrace;
    const double *hm, *xav, *yav, *xxav, *yyav;
    const double *mm;                  // ncat+1 planes, category 0 (open water) first
    const double *tm;                  // planes [n-1][nt], n = 1..ncat
    double *mc, *mx, *my;              // ncat+1 planes each
    double *tc, *tx, *ty;              // as tm
    signed char ttype[kRemapMaxTrace];          // tracer_type
    unsigned char parent[kRemapMaxTrace];       // depend - 1 (types 2, 3)
    signed char slot[kRemapMaxTrace];           // type 1 with dependents: index into the thread's mtxav / mtyav; else -1
};

namespace recon {
constexpr double puny = 1.0e-11;
__device__ __forceinline__ double min2(double a, double x) { return x < a ? x : a; }
__device__ __forceinline__ double max2(double a, double x) { return x > a ? x : a; }

// limited_gradient for one cell (:1404-1482).  ph[k], mk[k]: phi and phimask of the 3 x 3 neighbourhood,
// k = (dj+1)*3 + (di+1); cx, cy: cnx, cny of the cell
__device__ __forceinline__ void limited_gradient(const double (&ph)[9], const double (&mk)[9], double cx, double cy,
                                                 double &gx, double &gy)
{
    const double p = ph[4];
    auto nb = [&](int k) { return mk[k] * ph[k] + (1.0 - mk[k]) * p; };
    const double nw = nb(6), n = nb(7), ne = nb(8), w = nb(3), e = nb(5), sw = nb(0), s = nb(1), se = nb(2);
    const double gxtmp = (e - w) * 0.5, gytmp = (n - s) * 0.5;
    double pmn = nw, pmx = nw;
    const double lst[8] = {n, ne, w, p, e, sw, s, se};
#pragma unroll
    for (int k = 0; k < 8; ++k) { pmn = min2(pmn, lst[k]); pmx = max2(pmx, lst[k]); }
    pmn = pmn - p;
    pmx = pmx - p;
    double w1 = (0.5 - cx) * gxtmp + (0.5 - cy) * gytmp;
    double w2 = (0.5 - cx) * gxtmp - (0.5 + cy) * gytmp;
    const double w3 = -(0.5 + cx) * gxtmp - (0.5 + cy) * gytmp;
    const double w4 = (0.5 - cy) * gytmp - (0.5 + cx) * gxtmp;
    const double qmn = min2(min2(min2(w1, w2), w3), w4), qmx = max2(max2(max2(w1, w2), w3), w4);
    w1 = fabs(qmn) > fabs(pmn) ? max2(0.0, pmn / qmn) : 1.0;
    w2 = fabs(qmx) > fabs(pmx) ? max2(0.0, pmx / qmx) : 1.0;
    w1 = min2(w1, w2);
    gx = w1 * gxtmp;
    gy = w1 * gytmp;
}

__device__ __forceinline__ void load9(const double *f, size_t q, int ld, double (&v)[9])
{
#pragma unroll
    for (int dj = -1; dj <= 1; ++dj)
#pragma unroll
        for (int di = -1; di <= 1; ++di) v[(dj + 1) * 3 + di + 1] = f[q + (ptrdiff_t)dj * ld + di];
}
}  // namespace recon

// grid: (ceil(nxb / 128), nyb, ncat + 1); blockIdx.z = category
__global__ void __launch_bounds__(128) reconstruct_kernel(const __grid_constant__ ReconArgs a)
{
    using namespace recon;
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y, n = blockIdx.z;
    if (c >= a.nxb) return;
    const size_t q = (size_t)r * a.ld + c;
    const bool interior = c >= 1 && c <= a.nxb - 2 && r >= 1 && r <= a.nyb - 2;
    const double *mm = a.mm + (size_t)n * a.pe;
    const double m = mm[q];
    // icells list of this category: every cell of the block for open water (:930-943), the physical cells for ice
    // (:999-1010)
    const bool incell = (n == 0 || interior) && m > puny;
    const double mc = incell ? m : 0.0;
    double mx = 0.0, my = 0.0;
    double ph[9], mk[9];
    if (interior) {
        load9(mm, q, a.ld, ph);
        if (a.hm[q] > puny) {                                              // mass gradient, limited over the land mask
            load9(a.hm, q, a.ld, mk);
            limited_gradient(ph, mk, a.xav[q], a.yav[q], mx, my);
        }
    }
    a.mc[(size_t)n * a.pe + q] = mc;
    a.mx[(size_t)n * a.pe + q] = mx;
    a.my[(size_t)n * a.pe + q] = my;
    if (n == 0 || a.ntrace == 0) return;

    const size_t t0 = (size_t)(n - 1) * a.ntrace * a.pe + q;               // this cell in tracer plane 0 of category n
    if (!incell) {                                                         // no ice here: tc = tx = ty = 0 (:1107-1117)
        for (int nt = 0; nt < a.ntrace; ++nt) {
            const size_t o = t0 + (size_t)nt * a.pe;
            a.tc[o] = 0.0; a.tx[o] = 0.0; a.ty[o] = 0.0;
        }
        return;
    }
    // mmask of the neighbourhood (:958-971: 1 where mm > puny, ghost cells included); ph still holds mm
#pragma unroll
    for (int k = 0; k < 9; ++k) mk[k] = ph[k] > puny ? 1.0 : 0.0;
    const double xav = a.xav[q], yav = a.yav[q], xxav = a.xxav[q], yyav = a.yyav[q];
    const double mxav = (mx * xxav + mc * xav) / m, myav = (my * yyav + mc * yav) / m;     // :1139-1146
    double mtxav[kRemapMaxParents], mtyav[kRemapMaxParents];

    for (int nt = 0; nt < a.ntrace; ++nt) {
        const double *tm = a.tm + (size_t)(n - 1) * a.ntrace * a.pe + (size_t)nt * a.pe;
        const size_t o = t0 + (size_t)nt * a.pe;
        const double t = tm[q];
        double tc = 0.0, tx = 0.0, ty = 0.0;
        const int type = a.ttype[nt];
        if (type == 1) {                                                   // :1150-1208
            load9(tm, q, a.ld, ph);
            limited_gradient(ph, mk, mxav, myav, tx, ty);
            tc = t - tx * mxav - ty * myav;
            const int sl = a.slot[nt];
            if (sl >= 0) {                                                 // centre of area * tracer for its dependents
                double ax = 0.0, ay = 0.0;
                if (fabs(t) > puny) {                                      // tmask (:977-990), mm > puny holds here
                    const double w1 = mc * tc, w2 = mc * tx + mx * tc, w3 = mc * ty + my * tc;
                    const double w7 = 1.0 / (m * t);
                    ax = (w1 * xav + w2 * xxav) * w7;
                    ay = (w1 * yav + w3 * yyav) * w7;
                }
                mtxav[sl] = ax; mtyav[sl] = ay;
            }
        } else if (type == 2) {                                            // :1210-1229
            const int p1 = a.parent[nt], sl = a.slot[p1];
            const double *tp = a.tm + (size_t)(n - 1) * a.ntrace * a.pe + (size_t)p1 * a.pe;
            double pk[9], tk[9];
            load9(tp, q, a.ld, pk);
#pragma unroll
            for (int k = 0; k < 9; ++k) tk[k] = (mk[k] > puny && fabs(pk[k]) > puny) ? 1.0 : 0.0;   // tmask of the parent
            if (tk[4] > puny) {
                load9(tm, q, a.ld, ph);
                limited_gradient(ph, tk, mtxav[sl], mtyav[sl], tx, ty);
            }
            tc = t - tx * mtxav[sl] - ty * mtyav[sl];
        } else if (type == 3) {                                            // :1231-1239
            tc = t;
        }
        a.tc[o] = tc; a.tx[o] = tx; a.ty[o] = ty;
    }
}

}  // namespace evp
