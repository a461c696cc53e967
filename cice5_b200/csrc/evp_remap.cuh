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
    signed char slot[kRemapMaxTrace];           // tracer with dependents: its slot in the per-thread parent storage; else -1
    unsigned char order[kRemapMaxTrace];        // tracer numbers sorted by type (stable): parents come before children
    short n1, n2;                               // number of type-1, type-2 tracers (the rest is type 3)
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
    const double z = 0.0 * p;
    if (bits == 0x1ffu) {                                                  // every neighbour present (pack ice)
#pragma unroll
        for (int k = 0; k < 9; ++k) nb[k] = ph[k] + z;
    } else {
#pragma unroll
        for (int k = 0; k < 9; ++k) nb[k] = (bits >> k) & 1u ? ph[k] + z : 0.0 * ph[k] + p;
    }
    limit(nb, p, cx, cy, gx, gy);
}
}  // namespace recon

// ---- the kernel: one thread per cell and category ------------------------------------------------------------------
// grid: (ceil(nxb / 32), ceil(nyb / 4), ncat + 1), block 32 x 4: a warp is 32 consecutive cells of a row (256-byte
// coalesced loads and stores), the four rows of a block share their neighbour rows in L1; blockIdx.z = category.
// Measured on the 0.1-degree grid, 5 categories x 11 tracers (profiles/r02_reconstruct.md): the work is ~3000 issued
// instructions per cell and category (no FMA contraction, FP64 min / max as compare + select), a warp brings only 256
// new bytes per tracer plane and then waits for them, so the planes ahead are requested from L2 early
// (prefetch.global.L2); staging the tiles with TMA into a shared-memory ring was built and measured slower (8.4 ms
// against 6.5 ms: the producer / consumer lockstep per plane and 24 instead of 32 resident warps cost more than the
// exposed loads).
#ifndef EVP_RC_DIRECT_BLOCKS
#define EVP_RC_DIRECT_BLOCKS 8
#endif
#ifndef EVP_RC_DIRECT_H
#define EVP_RC_DIRECT_H 4
#endif
#ifndef EVP_RC_L2_AHEAD
#define EVP_RC_L2_AHEAD 4
#endif
constexpr int kRcDirectH = EVP_RC_DIRECT_H, kRcDirectW = 128 / kRcDirectH;
// dynamic shared memory: per tracer with dependents (slot) and thread, the centre of area * tracer (x, y) and the tmask
// bits of the neighbourhood (:977-990), kept for the type-2 tracers that follow: 18 bytes x 128 x slots
__global__ void __launch_bounds__(128, EVP_RC_DIRECT_BLOCKS) reconstruct_kernel(const __grid_constant__ ReconArgs a, int nparents)
{
    using namespace recon;
    extern __shared__ double par[];
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    double *mtxav = par + tid, *mtyav = par + (size_t)nparents * 128 + tid;
    unsigned short *tbits = reinterpret_cast<unsigned short *>(par + (size_t)2 * nparents * 128) + tid;
    // a block is a tile of kRcDirectW x kRcDirectH cells: the rows of a tile share their neighbour rows in L1
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y * blockDim.y + threadIdx.y, n = blockIdx.z;
    if (c >= a.nxb || r >= a.nyb) return;
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
    const double *tm0 = a.tm + t0;
#if EVP_RC_L2_AHEAD
    // A warp brings 256 new bytes per tracer plane and waits for them: too little in flight to keep HBM busy.  Ask L2
    // for this cell's line of the planes ahead now (the neighbour rows are the neighbour warps' lines)
#pragma unroll 1
    for (int nt = 1; nt < a.ntrace && nt <= EVP_RC_L2_AHEAD; ++nt)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(tm0 + (size_t)nt * a.pe));
#endif
    for (int nt = 0; nt < a.ntrace; ++nt) {
        const size_t o = t0 + (size_t)nt * a.pe;
        load9(tm0 + (size_t)nt * a.pe, 0, a.ld, ph);
#if EVP_RC_L2_AHEAD
        if (nt + EVP_RC_L2_AHEAD + 1 < a.ntrace)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(tm0 + (size_t)(nt + EVP_RC_L2_AHEAD + 1) * a.pe));
#endif
        const double t = ph[4];
        const int type = a.ttype[nt];
        double tc, tx = 0.0, ty = 0.0;
        if (type == 3) {                                                   // :1231-1239
            tc = t;
        } else {
            unsigned bits = mbits;                                         // type 1 (:1150-1208): mmask, centre of mass
            double cx = mxav, cy = myav;
            if (type == 2) {                                               // :1210-1229: the parent's tmask and centre
                const int sl = a.slot[a.parent[nt]] * 128;
                bits = tbits[sl]; cx = mtxav[sl]; cy = mtyav[sl];
            }
            if ((bits >> 4) & 1u) limited_gradient_bits(ph, bits, cx, cy, tx, ty);
            tc = t - tx * cx - ty * cy;
            const int sl = a.slot[nt];
            if (type == 1 && sl >= 0) {                                    // centre of area * tracer for its dependents
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
                mtxav[sl * 128] = ax; mtyav[sl * 128] = ay; tbits[sl * 128] = (unsigned short)tb;
            }
        }
        a.tc[o] = tc; a.tx[o] = tx; a.ty[o] = ty;
    }
}

// ---- third stage: locate_triangles + triangle_coordinates + transport_integrals, then update_fields ----------------
// reference: source/ice_transport_remap.F90:1680-3047 (locate_triangles), :3078-3187 (triangle_coordinates), :3199-3509
// (transport_integrals), :3517-3729 (update_fields), called by horizontal_remap :640-850 with the compile-time settings
// l_fixed_area = .false., integral_order = 3 (source/ice_transport_driver.F90:31,47); other settings are not built.
//
// The reference writes the triangles of a whole block to memory (vertices, areas, source cells, six index lists) and
// then sweeps the lists once per category and tracer.  Here one thread owns one cell edge: it works out the edge's six
// triangles, keeps the ten points they are made of, and per triangle that survives the area threshold integrates every
// category and tracer over it straight away; triangles never reach memory.  The flux of a (category, tracer) at an edge
// is the reference's sum over the groups in ascending order, accumulated in its output cell (which this thread alone
// touches).  The case analysis of locate_triangles is data here: one table row per triangle the reference assigns.
namespace flux {
// points: 0 CL  1 CR  2 DL  3 DR  4 IL  5 IR  6 IC (= ICL = ICR without l_fixed_area)  7 DM  8, 9 DL, DR after the
// departure points in the side cells have been moved to the cell boundary (:2238-2255)
enum { CL, CR, DL, DR, IL, IR, IC, DM, DL2, DR2 };
enum { TL, BL, TR, BR, TC, BC };                       // source cell: ishift_* / jshift_* (:1820-1898)
enum { PL, ML, PR, MR, PC, MC };                       // areafact: +/- areafac_l, _r, _c
struct Row { unsigned char ng, p1, p2, p3, sh, fac; };  // ng = 0: no triangle
// TL / BL triangles (:2034-2108), in the order of the if / elseif chain; the rows of a case in assignment order
__device__ const Row kLeft[4][2] = {
    {{1, CL, IL, DL, TL, ML}, {0, 0, 0, 0, 0, 0}},
    {{1, CL, DL, IL, BL, PL}, {0, 0, 0, 0, 0, 0}},
    {{1, CL, DL, IC, TL, PL}, {3, CL, IC, IL, BL, PL}},
    {{3, CL, IL, IC, TL, ML}, {1, CL, IC, DL, BL, ML}}};
// TR / BR triangles (:2114-2190)
__device__ const Row kRight[4][2] = {
    {{2, CR, DR, IR, TR, MR}, {0, 0, 0, 0, 0, 0}},
    {{2, CR, IR, DR, BR, PR}, {0, 0, 0, 0, 0, 0}},
    {{2, CR, IC, DR, TR, PR}, {3, CR, IR, IC, BR, PR}},
    {{3, CR, IC, IR, TR, MR}, {2, CR, DR, IC, BR, MR}}};
// TC / BC triangles (:2394-2923), groups 4, 5, 6
__device__ const Row kCentre[12][3] = {
    {{4, CL, CR, DL2, TC, MC}, {5, CR, DR2, DL2, TC, MC}, {6, DL2, DR2, DM, TC, MC}},
    {{4, CL, IC, DL2, TC, MC}, {5, CR, DR2, IC, TC, MC}, {6, IC, IC, DM, BC, PC}},
    {{4, CL, DL2, CR, BC, PC}, {5, CR, DL2, DR2, BC, PC}, {6, DL2, DM, DR2, BC, PC}},
    {{4, CL, DL2, IC, BC, PC}, {5, CR, IC, DR2, BC, PC}, {6, IC, IC, DM, TC, MC}},
    {{4, CL, IC, DL2, TC, MC}, {5, CR, IC, DR2, BC, PR}, {6, DL2, IC, DM, TC, MC}},
    {{4, CL, IC, DL2, TC, MC}, {5, CR, IC, DR2, BC, PR}, {6, IC, IC, DM, BC, PC}},
    {{4, CL, IC, DL2, TC, ML}, {5, CR, IC, DR2, BC, PC}, {6, DR2, IC, DM, BC, PC}},
    {{4, CL, IC, DL2, TC, ML}, {5, CR, IC, DR2, BC, PC}, {6, IC, IC, DM, TC, MC}},
    {{4, CL, DL2, IC, BC, PL}, {5, CR, DR2, IC, TC, MC}, {6, IC, DR2, DM, TC, MC}},
    {{4, CL, DL2, IC, BC, PL}, {5, CR, DR2, IC, TC, MC}, {6, IC, IC, DM, BC, PC}},
    {{4, CL, DL2, IC, BC, PC}, {5, CR, DR2, IC, TC, MR}, {6, IC, DL2, DM, BC, PC}},
    {{4, CL, DL2, IC, BC, PC}, {5, CR, DR2, IC, TC, MR}, {6, IC, IC, DM, TC, MC}}};
// (ishift, jshift) of TL BL TR BR TC BC for the east edge (row 0) and the north edge (row 1)
__device__ const signed char kShift[2][6][2] = {{{1, 1}, {0, 1}, {1, -1}, {0, -1}, {1, 0}, {0, 0}},
                                                {{-1, 1}, {-1, 0}, {1, 1}, {1, 0}, {0, 1}, {0, 0}}};
constexpr double puny = 1.0e-11, eps16 = 1.0e-16;
}  // namespace flux
#ifndef EVP_FLUX_CHUNK
#define EVP_FLUX_CHUNK 4
#endif
constexpr int kFluxChunk = EVP_FLUX_CHUNK;


struct FluxArgs {
    ReconArgs a;                       // sizes, tracer tables, the reconstructed fields mc mx my tc tx ty
    const double *dpx, *dpy, *dxu, *dyu;
    double *mflx[2], *mtflx[2];        // [east, north]: ncat+1 planes; ncat * ntrace planes
};

// grid: (2 * ceil(nxb / 128), nyb, ncat + 1); blockIdx.x = 2 * (column block) + edge (0 = east, 1 = north), blockIdx.z
// = category: a thread integrates one category (its mass and every tracer) over the triangles of one cell edge; the
// triangles are worked out again per category (a few hundred operations against the ~100 gathers of a category).  The
// east and the north edges of the same cells run side by side: they gather from the same 3 x 3 cells, once from HBM.
// Dynamic shared memory, per thread: the flux accumulators of the category (1 + ntrace doubles) and, per tracer with
// dependents, mtsum, mtxsum, mtysum of the current triangle (3 doubles x slots)
#ifndef EVP_FLUX_BLOCKS
#define EVP_FLUX_BLOCKS 8
#endif
__global__ void __launch_bounds__(128, EVP_FLUX_BLOCKS) edge_flux_kernel(const __grid_constant__ FluxArgs f, int nslots)
{
    using namespace flux;
    extern __shared__ double par[];
    const ReconArgs &a = f.a;
    const int c = (blockIdx.x >> 1) * blockDim.x + threadIdx.x, r = blockIdx.y;
    const int north = blockIdx.x & 1, n = blockIdx.z;
    if (c >= a.nxb) return;
    const size_t q = (size_t)r * a.ld + c;
    const int ntr = n > 0 ? a.ntrace : 0;
    double *acc = par + threadIdx.x;                                        // acc[k * 128]: k = 0 mass, 1 + nt tracers
    double *mtsum = par + (size_t)(1 + a.ntrace) * 128 + threadIdx.x, *mtxsum = mtsum + (size_t)nslots * 128,
           *mtysum = mtxsum + (size_t)nslots * 128;
    // the ten points, picked by table index: a per-thread array in local memory (in shared memory they cost occupancy:
    // measured slower, profiles/r02_remap.md)
    constexpr int kPS = 1;
    double pxs[10], pys[10];
    for (int k = 0; k <= ntr; ++k) acc[k * 128] = 0.0;                      // mflx = c0, mtflx = c0 (:3224-3230)
    // edges of the block that carry a flux (:1820-1826, :1860-1866); every other cell of the flux planes is zero
    const bool edge = north ? (c >= 1 && c <= a.nxb - 2 && r <= a.nyb - 2) : (c <= a.nxb - 2 && r >= 1 && r <= a.nyb - 2);
    const size_t ql = north ? q - 1 : q, qr = north ? q : q - a.ld;       // corners: left (i-1,j | i,j), right (i,j | i,j-1)
    bool moving = false;
    if (edge) moving = f.dpx[ql] != 0.0 || f.dpy[ql] != 0.0 || f.dpx[qr] != 0.0 || f.dpy[qr] != 0.0;        // :1956-1980
    if (moving) {
        const double fl = f.dxu[ql] * f.dyu[ql], fr = f.dxu[qr] * f.dyu[qr], fc = 0.5 * (fl + fr);           // areafac
        const double dxl = f.dpx[ql] / f.dxu[ql], dyl = f.dpy[ql] / f.dyu[ql];                               // :1986-1991
        const double dxr = f.dpx[qr] / f.dxu[qr], dyr = f.dpy[qr] / f.dyu[qr];
        pxs[CL * kPS] = -0.5; pys[CL * kPS] = 0.0; pxs[CR * kPS] = 0.5; pys[CR * kPS] = 0.0;
        double xdl, ydl, xdr, ydr;
        if (north) { xdl = pxs[CL * kPS] + dxl; ydl = pys[CL * kPS] + dyl; xdr = pxs[CR * kPS] + dxr; ydr = pys[CR * kPS] + dyr; }
        else { xdl = pxs[CL * kPS] - dyl; ydl = pys[CL * kPS] + dxl; xdr = pxs[CR * kPS] - dyr; ydr = pys[CR * kPS] + dxr; }           // rotated by pi/2
        const double xdm = 0.5 * (xdr + xdl), ydm = 0.5 * (ydr + ydl);
        const double yil = (pxs[CL * kPS] * (ydm - ydl) + xdm * ydl - xdl * ydm) / (xdm - xdl);
        const double yir = (pxs[CR * kPS] * (ydr - ydm) - xdm * ydr + xdr * ydm) / (xdr - xdm);
        const double md = (ydr - ydl) / (xdr - xdl);
        const double xic = fabs(md) > puny ? xdl - ydl / md : 0.0;
        pxs[DL * kPS] = xdl; pys[DL * kPS] = ydl; pxs[DR * kPS] = xdr; pys[DR * kPS] = ydr; pxs[IL * kPS] = pxs[CL * kPS]; pys[IL * kPS] = yil; pxs[IR * kPS] = pxs[CR * kPS]; pys[IR * kPS] = yir;
        pxs[IC * kPS] = xic; pys[IC * kPS] = 0.0; pxs[DM * kPS] = xdm; pys[DM * kPS] = ydm;
        const Row *rows[3] = {nullptr, nullptr, nullptr};
        int nrows[3] = {0, 0, 0};
        {
            int k = -1;
            if (yil > 0.0 && xdl < pxs[CL * kPS] && ydl >= 0.0) k = 0;
            else if (yil < 0.0 && xdl < pxs[CL * kPS] && ydl < 0.0) k = 1;
            else if (yil < 0.0 && xdl < pxs[CL * kPS] && ydl >= 0.0) k = 2;
            else if (yil > 0.0 && xdl < pxs[CL * kPS] && ydl < 0.0) k = 3;
            if (k >= 0) { rows[0] = kLeft[k]; nrows[0] = 2; }
            k = -1;
            if (yir > 0.0 && xdr >= pxs[CR * kPS] && ydr >= 0.0) k = 0;
            else if (yir < 0.0 && xdr >= pxs[CR * kPS] && ydr < 0.0) k = 1;
            else if (yir < 0.0 && xdr >= pxs[CR * kPS] && ydr >= 0.0) k = 2;
            else if (yir > 0.0 && xdr >= pxs[CR * kPS] && ydr < 0.0) k = 3;
            if (k >= 0) { rows[1] = kRight[k]; nrows[1] = 2; }
            if (xdl < pxs[CL * kPS]) { xdl = pxs[IL * kPS]; ydl = yil; }
            if (xdr > pxs[CR * kPS]) { xdr = pxs[IR * kPS]; ydr = yir; }
            pxs[DL2 * kPS] = xdl; pys[DL2 * kPS] = ydl; pxs[DR2 * kPS] = xdr; pys[DR2 * kPS] = ydr;
            k = -1;
            if (ydl >= 0.0 && ydr >= 0.0 && ydm >= 0.0) k = 0;
            else if (ydl >= 0.0 && ydr >= 0.0 && ydm < 0.0) k = 1;
            else if (ydl < 0.0 && ydr < 0.0 && ydm < 0.0) k = 2;
            else if (ydl < 0.0 && ydr < 0.0 && ydm >= 0.0) k = 3;
            else if (ydl >= 0.0 && ydr < 0.0 && xic >= 0.0 && ydm >= 0.0) k = 4;
            else if (ydl >= 0.0 && ydr < 0.0 && xic >= 0.0 && ydm < 0.0) k = 5;
            else if (ydl >= 0.0 && ydr < 0.0 && xic < 0.0 && ydm < 0.0) k = 6;
            else if (ydl >= 0.0 && ydr < 0.0 && xic < 0.0 && ydm >= 0.0) k = 7;
            else if (ydl < 0.0 && ydr >= 0.0 && xic < 0.0 && ydm >= 0.0) k = 8;
            else if (ydl < 0.0 && ydr >= 0.0 && xic < 0.0 && ydm < 0.0) k = 9;
            else if (ydl < 0.0 && ydr >= 0.0 && xic >= 0.0 && ydm < 0.0) k = 10;
            else if (ydl < 0.0 && ydr >= 0.0 && xic >= 0.0 && ydm >= 0.0) k = 11;
            if (k >= 0) { rows[2] = kCentre[k]; nrows[2] = 3; }
        }
        // the triangle of each group: the LAST row that names the group (a right-hand case overwrites group 3 of a
        // left-hand one, as the reference's assignments do)
        Row tri[6];
#pragma unroll
        for (int g = 0; g < 6; ++g) tri[g].ng = 0;
        for (int s3 = 0; s3 < 3; ++s3)
            for (int k = 0; k < nrows[s3]; ++k) {
                const Row w = rows[s3][k];
                if (w.ng) tri[w.ng - 1] = w;
            }
        const double *mcp = a.mc + (size_t)n * a.pe, *mxp = a.mx + (size_t)n * a.pe, *myp = a.my + (size_t)n * a.pe;
        const size_t t0 = (size_t)(n > 0 ? n - 1 : 0) * a.ntrace * a.pe;
        for (int g = 0; g < 6; ++g) {                                      // groups in ascending order (:3235)
            const Row w = tri[g];
            if (!w.ng) continue;
            double x1 = pxs[w.p1 * kPS], y1 = pys[w.p1 * kPS], x2 = pxs[w.p2 * kPS], y2 = pys[w.p2 * kPS], x3 = pxs[w.p3 * kPS], y3 = pys[w.p3 * kPS];
            const double fa = w.fac == PL ? fl : w.fac == ML ? -fl : w.fac == PR ? fr : w.fac == MR ? -fr : w.fac == PC ? fc : -fc;
            const double area = 0.5 * ((x2 - x1) * (y3 - y1) - (y2 - y1) * (x3 - x1)) * fa;                  // :2932-2937
            if (fabs(area) < eps16 * fc) continue;                          // not in the list of this group (:2939-2946)
            const int is = kShift[north][w.sh][0], js = kShift[north][w.sh][1];
            // vertices in the coordinates of the source cell (:2989-3021), then the four integration points (:3160-3183)
            if (north) {
                x1 = x1 - 1.0 * is; y1 = y1 + 0.5 - 1.0 * js; x2 = x2 - 1.0 * is; y2 = y2 + 0.5 - 1.0 * js;
                x3 = x3 - 1.0 * is; y3 = y3 + 0.5 - 1.0 * js;
            } else {
                double t;
                t = x1; x1 = y1 + 0.5 - 1.0 * is; y1 = -t - 1.0 * js;
                t = x2; x2 = y2 + 0.5 - 1.0 * is; y2 = -t - 1.0 * js;
                t = x3; x3 = y3 + 0.5 - 1.0 * is; y3 = -t - 1.0 * js;
            }
            const double p333 = 1.0 / 3.0;
            const double x0 = p333 * (x1 + x2 + x3), y0 = p333 * (y1 + y2 + y3);
            x1 = 0.4 * x1 + 0.6 * x0; y1 = 0.4 * y1 + 0.6 * y0;
            x2 = 0.4 * x2 + 0.6 * x0; y2 = 0.4 * y2 + 0.6 * y0;
            x3 = 0.4 * x3 + 0.6 * x0; y3 = 0.4 * y3 + 0.6 * y0;
            const size_t qs = q + (ptrdiff_t)js * a.ld + is;                // source cell (iflux, jflux)
            // :3380-3420, cubic (4-point) formula
            const double mc = mcp[qs], mx = mxp[qs], my = myp[qs];
            const double m0 = (-9.0 / 16.0) * (mc + x0 * mx + y0 * my);
            const double m1 = (25.0 / 48.0) * (mc + x1 * mx + y1 * my);
            const double m2 = (25.0 / 48.0) * (mc + x2 * mx + y2 * my);
            const double m3 = (25.0 / 48.0) * (mc + x3 * mx + y3 * my);
            const double msum = m0 + m1 + m2 + m3;
            acc[0] = acc[0] + area * msum;
            if (ntr == 0) continue;
            double w0 = m0 * x0, w1 = m1 * x1, w2 = m2 * x2, w3 = m3 * x3;
            const double mxsum = w0 + w1 + w2 + w3;
            const double mxxsum = w0 * x0 + w1 * x1 + w2 * x2 + w3 * x3;
            const double mxysum = w0 * y0 + w1 * y1 + w2 * y2 + w3 * y3;
            w0 = m0 * y0; w1 = m1 * y1; w2 = m2 * y2; w3 = m3 * y3;
            const double mysum = w0 + w1 + w2 + w3;
            const double myysum = w0 * y0 + w1 * y1 + w2 * y2 + w3 * y3;
            // :3423-3500, the tracers type by type (a tracer's flux depends on its parent's sums only, and parents are of
            // the type before).  A thread's loads are its latency: the fields of kFluxChunk tracers are requested
            // together before the first is used
            for (int nb = 0; nb < a.n1; nb += kFluxChunk) {                // type 1: mass * tracer
                double tcv[kFluxChunk], txv[kFluxChunk], tyv[kFluxChunk];
                int ntv[kFluxChunk];
#pragma unroll
                for (int k = 0; k < kFluxChunk; ++k) {
                    ntv[k] = a.order[nb + k < a.n1 ? nb + k : nb];
                    const size_t pl = t0 + (size_t)ntv[k] * a.pe + qs;
                    tcv[k] = a.tc[pl]; txv[k] = a.tx[pl]; tyv[k] = a.ty[pl];
                }
#pragma unroll
                for (int k = 0; k < kFluxChunk; ++k) {
                    if (nb + k >= a.n1) break;
                    const int nt = ntv[k], sl = a.slot[nt] * 128;
                    const double tc = tcv[k], tx = txv[k], ty = tyv[k];
                    const double ms = msum * tc + mxsum * tx + mysum * ty;
                    if (sl >= 0) {
                        mtsum[sl] = ms;
                        mtxsum[sl] = mxsum * tc + mxxsum * tx + mxysum * ty;
                        mtysum[sl] = mysum * tc + mxysum * tx + myysum * ty;
                    }
                    acc[(1 + nt) * 128] = acc[(1 + nt) * 128] + area * ms;
                }
            }
            const int e2 = a.n1 + a.n2;
            for (int nb = a.n1; nb < e2; nb += kFluxChunk) {               // type 2: mass * parent * tracer
                double tcv[kFluxChunk], txv[kFluxChunk], tyv[kFluxChunk];
                int ntv[kFluxChunk];
#pragma unroll
                for (int k = 0; k < kFluxChunk; ++k) {
                    ntv[k] = a.order[nb + k < e2 ? nb + k : nb];
                    const size_t pl = t0 + (size_t)ntv[k] * a.pe + qs;
                    tcv[k] = a.tc[pl]; txv[k] = a.tx[pl]; tyv[k] = a.ty[pl];
                }
#pragma unroll
                for (int k = 0; k < kFluxChunk; ++k) {
                    if (nb + k >= e2) break;
                    const int nt = ntv[k], sl = a.slot[nt] * 128, s1 = a.slot[a.parent[nt]] * 128;
                    const double ms = mtsum[s1] * tcv[k] + mtxsum[s1] * txv[k] + mtysum[s1] * tyv[k];
                    if (sl >= 0) mtsum[sl] = ms;
                    acc[(1 + nt) * 128] = acc[(1 + nt) * 128] + area * ms;
                }
            }
            for (int nb = e2; nb < ntr; ++nb) {                            // type 3: mass * grandparent * parent * tracer
                const int nt = a.order[nb], s1 = a.slot[a.parent[nt]] * 128;
                const double ms = mtsum[s1] * a.tc[t0 + (size_t)nt * a.pe + qs];
                acc[(1 + nt) * 128] = acc[(1 + nt) * 128] + area * ms;
            }
        }
    }
    f.mflx[north][(size_t)n * a.pe + q] = acc[0];
    const size_t t0 = (size_t)(n > 0 ? n - 1 : 0) * a.ntrace * a.pe;
    for (int nt = 0; nt < ntr; ++nt) f.mtflx[north][t0 + (size_t)nt * a.pe + q] = acc[(1 + nt) * 128];
}

struct UpdateArgs {
    ReconArgs a;
    const double *tarear;
    const double *mflx[2], *mtflx[2];
    double *mm, *tm;                   // updated in place on the physical cells (the reconstruction's input planes)
    int *stop;                         // [ncat+1]: 0 or 1 + r * nxb + c of the LAST cell (loop order) whose new area < -puny
};

// grid: (ceil(nxb / 128), nyb, ncat + 1).  Dynamic shared memory: per tracer with dependents and thread its old and its
// new value (2 x 8 bytes x 128 x slots)
__global__ void __launch_bounds__(128) update_fields_kernel(const __grid_constant__ UpdateArgs u, int nslots)
{
    extern __shared__ double par[];
    const ReconArgs &a = u.a;
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y, n = blockIdx.z;
    if (c < 1 || c > a.nxb - 2 || r < 1 || r > a.nyb - 2) return;
    const size_t q = (size_t)r * a.ld + c;
    double *told = par + threadIdx.x, *tnew = par + (size_t)nslots * 128 + threadIdx.x;
    const double *fe = u.mflx[0] + (size_t)n * a.pe, *fn = u.mflx[1] + (size_t)n * a.pe;
    const double tarear = u.tarear[q];
    const double mold = u.mm[(size_t)n * a.pe + q];
    double w1 = fe[q] - fe[q - 1] + fn[q] - fn[q - a.ld];                  // :3605-3609
    double m = mold - w1 * tarear;
    if (m < -1.0e-11) atomicMax(u.stop + n, 1 + r * a.nxb + c);            // the reference aborts (:3611-3630)
    else if (m < 0.0) m = 0.0;
    u.mm[(size_t)n * a.pe + q] = m;
    if (n == 0 || a.ntrace == 0) return;
    const size_t t0 = (size_t)(n - 1) * a.ntrace * a.pe;
    for (int nb = 0; nb < a.ntrace; nb += kFluxChunk) {                    // loads of kFluxChunk tracers in flight together
        double tv[kFluxChunk], wv[kFluxChunk];
#pragma unroll
        for (int k = 0; k < kFluxChunk; ++k) {
            const size_t pl = t0 + (size_t)(nb + k < a.ntrace ? nb + k : nb) * a.pe;
            const double *te = u.mtflx[0] + pl, *tn_ = u.mtflx[1] + pl;
            tv[k] = u.tm[pl + q];
            wv[k] = te[q] - te[q - 1] + tn_[q] - tn_[q - a.ld];            // used only where the reference evaluates it
        }
#pragma unroll
        for (int k = 0; k < kFluxChunk; ++k) {
            const int nt = nb + k;
            if (nt >= a.ntrace) break;
            const size_t pl = t0 + (size_t)nt * a.pe;
            const int type = a.ttype[nt], sl = a.slot[nt] * 128;
            const double t = tv[k];
            double mtold, den = m, tn = 0.0;                               // :3560-3590, :3645-3700
            bool on = m > 0.0;
            if (type == 1) {
                mtold = mold * t;
            } else {
                const int s1 = a.slot[a.parent[nt]] * 128;
                if (type == 2) {
                    mtold = mold * told[s1] * t;
                    on = on && fabs(tnew[s1]) > 0.0;
                    den = m * tnew[s1];
                } else {
                    const int s2 = a.slot[a.parent[a.parent[nt]]] * 128;
                    mtold = mold * told[s2] * told[s1] * t;
                    on = on && fabs(tnew[s1]) > 0.0 && fabs(tnew[s2]) > 0.0;
                    den = m * tnew[s2] * tnew[s1];
                }
            }
            if (on) tn = (mtold - wv[k] * tarear) / den;
            if (sl >= 0) { told[sl] = t; tnew[sl] = tn; }
            u.tm[pl + q] = tn;
        }
    }
}

// ---- the caller's conversions: state_to_tracers (source/ice_transport_driver.F90:789-900) and tracers_to_state
// (:908-1002) of transport_remap ------------------------------------------------------------------------------------------
struct StateArgs {
    int nxb, nyb, ld;
    size_t pe;
    int ncat, ntrcr, nt_qsno, nslyr;   // nt_qsno 1-based (ice_state)
    double rhosLf;                     // rhos * Lfresh
    double *aice0, *aicen, *vicen, *vsnon, *trcrn;     // 1, ncat, ncat, ncat, ncat * ntrcr planes ([n-1][it-1])
    double *aim, *trm;                 // ncat + 1, ncat * (2 + ntrcr) planes: mm, tm of the reconstruction
};

// grid: (ceil(nxb / 128), nyb, ncat + 1): every cell of the block, ghost cells included
__global__ void state_to_tracers_kernel(const __grid_constant__ StateArgs s)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y, n = blockIdx.z;
    if (c >= s.nxb) return;
    const size_t q = (size_t)r * s.ld + c;
    if (n == 0) { s.aim[q] = s.aice0[q]; return; }
    const int ntrace = 2 + s.ntrcr;
    const double am = s.aicen[(size_t)(n - 1) * s.pe + q];
    s.aim[(size_t)n * s.pe + q] = am;
    double *tr = s.trm + (size_t)(n - 1) * ntrace * s.pe + q;
    const double *tn = s.trcrn + (size_t)(n - 1) * s.ntrcr * s.pe + q;
    if (am > 1.0e-11) {
        const double w1 = 1.0 / am;
        tr[0] = s.vicen[(size_t)(n - 1) * s.pe + q] * w1;                  // hice
        tr[s.pe] = s.vsnon[(size_t)(n - 1) * s.pe + q] * w1;               // hsno
        for (int it = 1; it <= s.ntrcr; ++it) {
            const bool snow = it >= s.nt_qsno && it < s.nt_qsno + s.nslyr; // snow enthalpy, shifted (:880-886)
            const double v = tn[(size_t)(it - 1) * s.pe];
            tr[(size_t)(1 + it) * s.pe] = snow ? v + s.rhosLf : v;
        }
    } else {
        for (int k = 0; k < ntrace; ++k) tr[(size_t)k * s.pe] = 0.0;
    }
}

// cells whose new area is not positive keep their old state (:963-975)
__global__ void tracers_to_state_kernel(const __grid_constant__ StateArgs s)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y, n = blockIdx.z;
    if (c >= s.nxb) return;
    const size_t q = (size_t)r * s.ld + c;
    if (n == 0) { s.aice0[q] = s.aim[q]; return; }
    const int ntrace = 2 + s.ntrcr;
    const double am = s.aim[(size_t)n * s.pe + q];
    if (!(am > 0.0)) return;
    const double *tr = s.trm + (size_t)(n - 1) * ntrace * s.pe + q;
    double *tn = s.trcrn + (size_t)(n - 1) * s.ntrcr * s.pe + q;
    s.aicen[(size_t)(n - 1) * s.pe + q] = am;
    s.vicen[(size_t)(n - 1) * s.pe + q] = am * tr[0];
    s.vsnon[(size_t)(n - 1) * s.pe + q] = am * tr[s.pe];
    for (int it = 1; it <= s.ntrcr; ++it) {
        const bool snow = it >= s.nt_qsno && it < s.nt_qsno + s.nslyr;
        const double v = tr[(size_t)(1 + it) * s.pe];
        tn[(size_t)(it - 1) * s.pe] = snow ? v - s.rhosLf : v;
    }
}

}  // namespace evp
