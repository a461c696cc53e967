// evp_remap.cuh -- first stages of horizontal transport on the device: departure_points and the
// field reconstruction (make_masks + construct_fields + limited_gradient, second half of this file).
//
// reference: source/ice_transport_remap.F90:1493-1670 (departure_points, nghost = 1), called by horizontal_remap
// (:541-550) right after evp in step_dynamics (source/ice_step_mod.F90:1126-1130); the ghost cells of dpx, dpy are then
// filled by two ice_HaloUpdate calls (NE corner, vector; :574-578) -- the generic halo of evp_prep.cuh.  Reads uvel,
// vvel straight from the resident EVP state: the velocities never leave the device between evp and this stage.
#pragma once
#include "evp_kernels.cuh"
#include "evp_subcycle_tma.cuh"     // mbarrier / TMA helpers

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

__device__ __forceinline__ void load9(const double *f, size_t q, int ld, double (&v)[9])
{
#pragma unroll
    for (int dj = -1; dj <= 1; ++dj)
#pragma unroll
        for (int di = -1; di <= 1; ++di) v[(dj + 1) * 3 + di + 1] = f[q + (ptrdiff_t)dj * ld + di];
}

// the limiter of limited_gradient (:1431-1482) on the eight masked neighbour values nb[k] (k = (dj+1)*3 + (di+1), nb[4]
// unused) and the cell's own value p
__device__ __forceinline__ void limit(const double (&nb)[9], double p, double cx, double cy, double &gx, double &gy)
{
    const double gxtmp = (nb[5] - nb[3]) * 0.5, gytmp = (nb[7] - nb[1]) * 0.5;
    double pmn = nb[6], pmx = nb[6];                                       // nw, n, ne, w, phi(i,j), e, sw, s, se
    const double lst[8] = {nb[7], nb[8], nb[3], p, nb[5], nb[0], nb[1], nb[2]};
#pragma unroll
    for (int k = 0; k < 8; ++k) { pmn = min2(pmn, lst[k]); pmx = max2(pmx, lst[k]); }
    pmn = pmn - p;
    pmx = pmx - p;
    double w1 = (0.5 - cx) * gxtmp + (0.5 - cy) * gytmp;
    double w2 = (0.5 - cx) * gxtmp - (0.5 + cy) * gytmp;
    const double w3 = -(0.5 + cx) * gxtmp - (0.5 + cy) * gytmp;
    const double w4 = (0.5 - cy) * gytmp - (0.5 + cx) * gxtmp;
    const double qmn = min2(min2(min2(w1, w2), w3), w4), qmx = max2(max2(max2(w1, w2), w3), w4);
    // max(c0, pmn/qmn), max(c0, pmx/qmx) (:1463-1474).  A cell that is the smallest (largest) of its neighbourhood has
    // pmn (pmx) = 0: the quotient is a zero of either sign and max(c0, .) returns c0 -- decided here without the
    // division, whose zero-numerator case leaves the fast path of the FP64 division sequence (in the first profile
    // that slow path, at 6 of 32 lanes, was 40 % of all issued instructions)
    w1 = 1.0;
    if (fabs(qmn) > fabs(pmn)) w1 = pmn == 0.0 ? 0.0 : max2(0.0, pmn / qmn);
    w2 = 1.0;
    if (fabs(qmx) > fabs(pmx)) w2 = pmx == 0.0 ? 0.0 : max2(0.0, pmx / qmx);
    w1 = min2(w1, w2);
    gx = w1 * gxtmp;
    gy = w1 * gytmp;
}

// limited_gradient for one cell with an arbitrary mask field (the mass gradient over hm): phimask * phi + (1 - phimask)
// * phi(i,j) as written (:1406-1421)
__device__ __forceinline__ void limited_gradient(const double (&ph)[9], const double (&mk)[9], double cx, double cy,
                                                 double &gx, double &gy)
{
    double nb[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) nb[k] = mk[k] * ph[k] + (1.0 - mk[k]) * ph[4];
    limit(nb, ph[4], cx, cy, gx, gy);
}

// the same with a mask known to be exactly 1 or 0 (mmask, tmask; bit k of `bits`): 1 * phi(nb) + 0 * phi(i,j) where the
// bit is set, 0 * phi(nb) + 1 * phi(i,j) where not -- the reference's two products and sum with the multiplications by
// one and the subtraction 1 - mask (exact) folded; the product with zero stays (it carries the sign of a zero sum)
__device__ __forceinline__ void limited_gradient_bits(const double (&ph)[9], unsigned bits, double cx, double cy,
                                                      double &gx, double &gy)
{
    double nb[9];
    const double p = ph[4];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const bool on = (bits >> k) & 1u;
        const double keep = on ? ph[k] : p, drop = on ? p : ph[k];
        nb[k] = on ? keep + 0.0 * drop : 0.0 * drop + keep;
    }
    limit(nb, p, cx, cy, gx, gy);
}
}  // namespace recon

// ---- the product kernel: tiles staged by TMA ---------------------------------------------------------------------
// A work item is one tile of kRcTX x kRcTY cells of one category.  Consumer warp w owns tile row w (one cell per
// thread); a producer warp walks the same item / plane sequence and brings the (kRcTX + 4) x (kRcTY + 2) box of each plane
// -- mm of the category, hm, then the tracer planes in table order -- into a ring of kRcStages shared-memory stages with
// cp.async.bulk.tensor, so that no consumer ever waits on a global load: it copies its 3 x 3 neighbourhood from the
// stage to registers, hands the stage back and does the FP64 work.  Boxes that stick out of the block are zero-filled
// by the TMA unit (those values are never used: only physical cells read their neighbours).  Persistent CTAs, items dealt
// round-robin with the category as the fastest index (CTAs that run side by side share hm and the halo rows in L2).
#ifndef EVP_RC_TY
#define EVP_RC_TY 8
#endif
#ifndef EVP_RC_STAGES
#define EVP_RC_STAGES 6
#endif
#ifndef EVP_RC_BLOCKS
#define EVP_RC_BLOCKS 3
#endif
constexpr int kRcTX = 32, kRcTY = EVP_RC_TY, kRcStages = EVP_RC_STAGES, kRcBlocksPerSM = EVP_RC_BLOCKS;
constexpr int kRcBoxW = kRcTX + 4, kRcBoxH = kRcTY + 2;      // box starts must be 16-byte aligned: two columns each side
constexpr int kRcBoxBytes = kRcBoxW * kRcBoxH * 8;                         // 2880
constexpr int kRcStageBytes = (kRcBoxBytes + 127) / 128 * 128;             // 2944

struct ReconTmaArgs {
    ReconArgs a;
    int tiles_x, tiles_y;
    int plane_hm, plane_mm, plane_tm;          // plane numbers (z coordinate of the tensor map) of hm, mm(0), tm(1,1)
};

__global__ void __launch_bounds__((kRcTY + 1) * 32, kRcBlocksPerSM)
reconstruct_tma_kernel(const __grid_constant__ CUtensorMap map, const __grid_constant__ ReconTmaArgs p)
{
    using namespace recon;
    __shared__ __align__(128) unsigned char ring[kRcStages * kRcStageBytes];
    __shared__ unsigned long long full_bar[kRcStages], empty_bar[kRcStages];
    const ReconArgs &a = p.a;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kRcStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], kRcTY); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    const int ncat1 = a.ncat + 1;
    const long long nitems = (long long)p.tiles_x * p.tiles_y * ncat1;
    int stage = 0;
    unsigned phase = 0;

    if (warp == kRcTY) {
        // ------------------------------ producer ------------------------------
        if (lane != 0) return;
        for (long long it = blockIdx.x; it < nitems; it += gridDim.x) {
            const int n = (int)(it % ncat1);
            const long long tile = it / ncat1;
            const int x0 = (int)(tile % p.tiles_x) * kRcTX - 2, y0 = (int)(tile / p.tiles_x) * kRcTY - 1;
            const int nplanes = 2 + (n > 0 ? a.ntrace : 0);
            for (int k = 0; k < nplanes; ++k) {
                const int z = k == 0 ? p.plane_mm + n : k == 1 ? p.plane_hm : p.plane_tm + (n - 1) * a.ntrace + (k - 2);
                mbar_wait(&empty_bar[stage], phase ^ 1u);
                mbar_expect_tx(&full_bar[stage], kRcBoxBytes);
                tma_load_3d(ring + (size_t)stage * kRcStageBytes, &map, &full_bar[stage], x0, y0, z);
                if (++stage == kRcStages) { stage = 0; phase ^= 1u; }
            }
        }
        return;
    }

    // ------------------------------ consumers ------------------------------
    // stage -> registers: the 3 x 3 neighbourhood of this thread's cell (or only its centre), then the stage goes back
    auto take = [&](double (&v)[9], bool all) {
        mbar_wait(&full_bar[stage], phase);
        const double *b = reinterpret_cast<const double *>(ring + (size_t)stage * kRcStageBytes) + (warp + 1) * kRcBoxW + lane + 2;
        if (all) {
#pragma unroll
            for (int dj = -1; dj <= 1; ++dj)
#pragma unroll
                for (int di = -1; di <= 1; ++di) v[(dj + 1) * 3 + di + 1] = b[dj * kRcBoxW + di];
        } else {
            v[4] = b[0];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[stage]);
        if (++stage == kRcStages) { stage = 0; phase ^= 1u; }
    };

    for (long long it = blockIdx.x; it < nitems; it += gridDim.x) {
        const int n = (int)(it % ncat1);
        const long long tile = it / ncat1;
        const int c = (int)(tile % p.tiles_x) * kRcTX + lane, r = (int)(tile / p.tiles_x) * kRcTY + warp;
        const bool valid = c < a.nxb && r < a.nyb;
        const bool interior = c >= 1 && c <= a.nxb - 2 && r >= 1 && r <= a.nyb - 2;
        const size_t q = (size_t)r * a.ld + c;
        double xav = 0.0, yav = 0.0, xxav = 0.0, yyav = 0.0;
        if (interior) { xav = a.xav[q]; yav = a.yav[q]; if (n > 0) { xxav = a.xxav[q]; yyav = a.yyav[q]; } }

        double ph[9];
        take(ph, interior);                                                // mm of this category
        const double m = ph[4];
        // icells list of this category: every cell of the block for open water (:930-943), the physical cells for ice
        // (:999-1010)
        const bool incell = valid && (n == 0 || interior) && m > puny;
        const double mc = incell ? m : 0.0;
        double mx = 0.0, my = 0.0;
        unsigned mbits = 0;                                                // mmask of the neighbourhood (:958-971)
        if (interior) {
#pragma unroll
            for (int k = 0; k < 9; ++k) mbits |= (ph[k] > puny ? 1u : 0u) << k;
        }
        {
            double mk[9];
            take(mk, interior);                                            // hm
            if (interior && mk[4] > puny) limited_gradient(ph, mk, xav, yav, mx, my);     // limited over the land mask
        }
        if (valid) {
            a.mc[(size_t)n * a.pe + q] = mc;
            a.mx[(size_t)n * a.pe + q] = mx;
            a.my[(size_t)n * a.pe + q] = my;
        }
        if (n == 0 || a.ntrace == 0) continue;

        const size_t t0 = (size_t)(n - 1) * a.ntrace * a.pe + q;           // this cell in tracer plane 0 of category n
        double mxav = 0.0, myav = 0.0;
        if (incell) {                                                      // :1139-1146
            const double nx = mx * xxav + mc * xav, ny = my * yyav + mc * yav;
            mxav = nx == 0.0 ? nx : nx / m;                                // m > 0: a zero keeps its sign; no slow path
            myav = ny == 0.0 ? ny : ny / m;
        }
        // per tracer with dependents: centre of area * tracer and its tmask over the neighbourhood (:977-990), kept for
        // the type-2 tracers that follow
        double mtxav[kRemapMaxParents], mtyav[kRemapMaxParents];
        unsigned short tbits[kRemapMaxParents];

        for (int nt = 0; nt < a.ntrace; ++nt) {
            const size_t o = t0 + (size_t)nt * a.pe;
            double tc = 0.0, tx = 0.0, ty = 0.0;
            const int type = a.ttype[nt];
            if (type == 1) {                                               // :1150-1208
                take(ph, incell);
                if (incell) {
                    const double t = ph[4];
                    limited_gradient_bits(ph, mbits, mxav, myav, tx, ty);
                    tc = t - tx * mxav - ty * myav;
                    const int sl = a.slot[nt];
                    if (sl >= 0) {                                         // centre of area * tracer for its dependents
                        unsigned tb = 0;
#pragma unroll
                        for (int k = 0; k < 9; ++k) tb |= (fabs(ph[k]) > puny ? 1u : 0u) << k;
                        tb &= mbits;
                        double ax = 0.0, ay = 0.0;
                        if ((tb >> 4) & 1u) {
                            const double w1 = mc * tc, w2 = mc * tx + mx * tc, w3 = mc * ty + my * tc;
                            const double w7 = 1.0 / (m * t);
                            ax = (w1 * xav + w2 * xxav) * w7;
                            ay = (w1 * yav + w3 * yyav) * w7;
                        }
                        mtxav[sl] = ax; mtyav[sl] = ay; tbits[sl] = (unsigned short)tb;
                    }
                }
            } else if (type == 2) {                                        // :1210-1229
                const int sl = a.slot[a.parent[nt]];
                const unsigned tb = incell ? tbits[sl] : 0u;
                const bool grad = (tb >> 4) & 1u;
                take(ph, grad);
                if (incell) {
                    const double ax = mtxav[sl], ay = mtyav[sl];
                    if (grad) limited_gradient_bits(ph, tb, ax, ay, tx, ty);
                    tc = ph[4] - tx * ax - ty * ay;
                }
            } else {                                                       // :1231-1239
                take(ph, false);
                if (incell) tc = ph[4];
            }
            if (valid) { a.tc[o] = tc; a.tx[o] = tx; a.ty[o] = ty; }
        }
    }
}

// ---- the same work, one thread per cell and category straight from global memory (EVP_B200_KERNEL=sweep: the build without TMA) ---------
// grid: (ceil(nxb / 128), nyb, ncat + 1); blockIdx.z = category
__global__ void __launch_bounds__(128, 8) reconstruct_kernel(const __grid_constant__ ReconArgs a)
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
    double ph[9];
    unsigned mbits = 0;                                                    // mmask of the neighbourhood (:958-971)
    if (interior) {
        load9(mm, q, a.ld, ph);
#pragma unroll
        for (int k = 0; k < 9; ++k) mbits |= (ph[k] > puny ? 1u : 0u) << k;
        if (a.hm[q] > puny) {                                              // mass gradient, limited over the land mask
            double mk[9];
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
    const double xav = a.xav[q], yav = a.yav[q], xxav = a.xxav[q], yyav = a.yyav[q];
    const double nx = mx * xxav + mc * xav, ny = my * yyav + mc * yav;                     // :1139-1146
    const double mxav = nx == 0.0 ? nx : nx / m, myav = ny == 0.0 ? ny : ny / m;           // m > 0: a zero keeps its sign
    // per tracer with dependents: centre of area * tracer and its tmask over the neighbourhood (:977-990), kept for the
    // type-2 tracers that follow
    double mtxav[kRemapMaxParents], mtyav[kRemapMaxParents];
    unsigned short tbits[kRemapMaxParents];

    for (int nt = 0; nt < a.ntrace; ++nt) {
        const double *tm = a.tm + (size_t)(n - 1) * a.ntrace * a.pe + (size_t)nt * a.pe;
        const size_t o = t0 + (size_t)nt * a.pe;
        double tc = 0.0, tx = 0.0, ty = 0.0;
        const int type = a.ttype[nt];
        if (type == 1) {                                                   // :1150-1208
            load9(tm, q, a.ld, ph);
            const double t = ph[4];
            limited_gradient_bits(ph, mbits, mxav, myav, tx, ty);
            tc = t - tx * mxav - ty * myav;
            const int sl = a.slot[nt];
            if (sl >= 0) {                                                 // centre of area * tracer for its dependents
                unsigned tb = 0;
#pragma unroll
                for (int k = 0; k < 9; ++k) tb |= (fabs(ph[k]) > puny ? 1u : 0u) << k;
                tb &= mbits;
                double ax = 0.0, ay = 0.0;
                if ((tb >> 4) & 1u) {
                    const double w1 = mc * tc, w2 = mc * tx + mx * tc, w3 = mc * ty + my * tc;
                    const double w7 = 1.0 / (m * t);
                    ax = (w1 * xav + w2 * xxav) * w7;
                    ay = (w1 * yav + w3 * yyav) * w7;
                }
                mtxav[sl] = ax; mtyav[sl] = ay; tbits[sl] = (unsigned short)tb;
            }
        } else if (type == 2) {                                            // :1210-1229
            const int sl = a.slot[a.parent[nt]];
            const unsigned tb = tbits[sl];
            const double ax = mtxav[sl], ay = mtyav[sl];
            double t;
            if ((tb >> 4) & 1u) {
                load9(tm, q, a.ld, ph);
                t = ph[4];
                limited_gradient_bits(ph, tb, ax, ay, tx, ty);
            } else {
                t = tm[q];
            }
            tc = t - tx * ax - ty * ay;
        } else if (type == 3) {                                            // :1231-1239
            tc = tm[q];
        }
        a.tc[o] = tc; a.tx[o] = tx; a.ty[o] = ty;
    }
}

}  // namespace evp
