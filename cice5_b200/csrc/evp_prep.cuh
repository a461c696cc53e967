// evp_prep.cuh -- once-per-call kernels around the subcycle loop and the generic width-1 halo.
//
// Reference semantics (path:line in the reference checkout):
//   evp_prep1   source/ice_dyn_shared.F90:270-365      to_ugrid  source/ice_grid.F90:1834-1876
//   evp_prep2   source/ice_dyn_shared.F90:377-614      to_tgrid  source/ice_grid.F90:1921-1958
//   evp_finish  source/ice_dyn_shared.F90:757-844
//   ice_HaloUpdate          serial/ice_boundary.F90:630-896 (+ MPI ring pre-fill mpi/..:2625-2632)
//   ice_HaloUpdate_stress   serial/ice_boundary.F90:3269-3443
// None of these is performance critical (a dozen field sweeps against 120 x 49 in the loop).
#pragma once
#include "evp_kernels.cuh"

namespace evp {

struct Geom {
    int nxl, ny, ld, nx_global, lo, nranks, rank;
    int tripole, ew_cyclic;
    int inbox_ld;
    int start[kMaxRanks + 1];      // x-slabs: rank t owns global columns start[t]+1 .. start[t+1]
};

__device__ __forceinline__ int glob_col(const Geom &g, int c)
{
    int ig = g.lo + c;
    if (ig < 1) ig += g.nx_global;
    if (ig > g.nx_global) ig -= g.nx_global;
    return ig;
}

// ---- inbox: all cross-rank halo traffic lands here (never directly in a peer's planes, so a
// producer kernel that sweeps whole planes cannot race with a neighbour that runs ahead).
// Two buffers, alternated per halo operation: a writer of operation n+2 has passed the sync of
// operation n+1, i.e. the owner has finished reading buffer (n & 1).
//   rows[4][inbox_ld]  : [field a: row ny-1 | row ny][field b: row ny-1 | row ny] of the fold source
//   cols[4][col_ld]    : [a: west ghost | a: east ghost | b: west ghost | b: east ghost], index = row
struct InboxGeom {
    int inbox_ld, col_ld;
};
__host__ __device__ __forceinline__ size_t inbox_doubles(const InboxGeom &ib)
{
    return (size_t)4 * ib.inbox_ld + (size_t)4 * ib.col_ld;
}
__device__ __forceinline__ double *inbox_row(double *base, const InboxGeom &ib, int row)
{
    return base + (size_t)row * ib.inbox_ld;
}
__device__ __forceinline__ double *inbox_col(double *base, const InboxGeom &ib, int col)
{
    return base + (size_t)4 * ib.inbox_ld + (size_t)col * ib.col_ld;
}

// ---- generic halo, step 1: my physical edge values go to whoever holds them as ghosts -------
// E-W: column 1 -> east-ghost inbox column of the west neighbour, column nxl -> west-ghost inbox
// column of the east neighbour (rows 1..ny).  Tripole: rows ny-1, ny of every physical column go
// to the inbox rows of the rank(s) whose columns fold onto it.  ioff = 0 for center fields
// (source nx+1-ig), 1 for NEcorner fields (source nx-ig) (serial/ice_boundary.F90:803-812,874-877).
struct PushArgs {
    Geom g;
    InboxGeom ib;
    const double *plane;
    double *inbox[kMaxRanks];         // buffer of this operation on every rank (peer pointers)
    int west, east;                   // neighbour ranks, -1 on a closed edge
    int ioff;
    int ew;                           // 0: skip the E-W columns (ice_HaloUpdate_stress)
};

__global__ void halo_push_kernel(const __grid_constant__ PushArgs a)
{
    const Geom &g = a.g;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < g.ny) {
        if (!a.ew) return;
        const int r = t + 1;
        if (a.west >= 0) inbox_col(a.inbox[a.west], a.ib, 1)[r] = a.plane[(size_t)r * g.ld + 1];
        if (a.east >= 0) inbox_col(a.inbox[a.east], a.ib, 0)[r] = a.plane[(size_t)r * g.ld + g.nxl];
    } else if (g.tripole && t < g.ny + g.nxl) {
        const int c = t - g.ny + 1;
        const int ig = g.lo + c;
        int X = g.nx_global + 1 - a.ioff - ig;
        if (X == 0) X = g.nx_global;
        int tr, ct, wt;
        slab_owner(g.start, g.nranks, X, tr, ct, wt);
#pragma unroll
        for (int rs = 0; rs < 2; ++rs) {
            const double v = a.plane[(size_t)(g.ny - 1 + rs) * g.ld + c];
            inbox_row(a.inbox[tr], a.ib, rs)[ct] = v;
            if (ct == 1) {
                const int t2 = (tr + g.nranks - 1) % g.nranks;
                inbox_row(a.inbox[t2], a.ib, rs)[g.start[t2 + 1] - g.start[t2] + 1] = v;
            }
            if (ct == wt) inbox_row(a.inbox[(tr + 1) % g.nranks], a.ib, rs)[0] = v;
        }
    }
}

// ---- generic halo, step 2: ring pre-fill, E-W ghosts and tripole fold out of my inbox -----------
// loc 0 = center (ghost row only), 1 = NEcorner (top physical row averaged + ghost row).
// sgn = +1 scalar, -1 vector.
__global__ void halo_apply_kernel(double *plane, Geom g, InboxGeom ib, double *inbox, int loc,
                                  double sgn, double fill, int west_closed, int east_closed, int first_pair)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int nxb = g.nxl + 2;
    if (t < nxb) {
        const int c = t;
        plane[c] = fill;                                           // south ghost row
        if (g.tripole) {
            const double *in_below = inbox_row(inbox, ib, 0);
            const double *in_top = inbox_row(inbox, ib, 1);
            if (loc == 0) {
                plane[(size_t)(g.ny + 1) * g.ld + c] = sgn * in_top[c];
            } else {
                double own;                                        // row ny before the fold
                if (c == 0) own = west_closed ? fill : inbox_col(inbox, ib, 0)[g.ny];
                else if (c == g.nxl + 1) own = east_closed ? fill : inbox_col(inbox, ib, 1)[g.ny];
                else own = plane[(size_t)g.ny * g.ld + c];
                fold_necorner_column(plane, g.ld, g.ny, c, glob_col(g, c), g.nx_global, own,
                                     in_below[c], in_top[c], sgn, first_pair);
            }
        } else {
            plane[(size_t)(g.ny + 1) * g.ld + c] = fill;           // open / closed north
        }
    } else if (t < nxb + g.ny) {
        const int r = t - nxb + 1;
        if (g.tripole && loc == 1 && r == g.ny) return;            // written by the fold above
        plane[(size_t)r * g.ld] = west_closed ? fill : inbox_col(inbox, ib, 0)[r];
        plane[(size_t)r * g.ld + g.nxl + 1] = east_closed ? fill : inbox_col(inbox, ib, 1)[r];
    }
}

// ---- the generic halo for a stack of planes in one go (transport: mc mx my tc tx ty of every category) ------------
// Plane p of the stack (blockIdx.y) uses field (p & 1) of slot (p >> 1) of a wide inbox -- kWideSlots slots with the
// layout above, two buffers alternated per operation like the narrow ones.  Same values into the same ghost cells as
// nplanes calls of halo_push_kernel / halo_apply_kernel; one launch each and, between ranks, one sync instead of nplanes.
// Scalar and vector fields share an operation: planes [neg_lo[0], neg_hi[0]) and [neg_lo[1], neg_hi[1]) of the stack change sign
// across the tripole fold (field_type_vector), the others do not.
constexpr int kWideSlots = 96;

__global__ void halo_push_multi_kernel(const __grid_constant__ PushArgs a, size_t plane_stride)
{
    const Geom &g = a.g;
    const int t = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
    const double *plane = a.plane + (size_t)p * plane_stride;
    const size_t so = (size_t)(p >> 1) * inbox_doubles(a.ib);
    const int fo = (p & 1) * 2;
    if (t < g.ny) {
        const int r = t + 1;
        if (a.west >= 0) inbox_col(a.inbox[a.west] + so, a.ib, fo + 1)[r] = plane[(size_t)r * g.ld + 1];
        if (a.east >= 0) inbox_col(a.inbox[a.east] + so, a.ib, fo)[r] = plane[(size_t)r * g.ld + g.nxl];
    } else if (g.tripole && t < g.ny + g.nxl) {
        const int c = t - g.ny + 1;
        const int ig = g.lo + c;
        int X = g.nx_global + 1 - a.ioff - ig;
        if (X == 0) X = g.nx_global;
        int tr, ct, wt;
        slab_owner(g.start, g.nranks, X, tr, ct, wt);
#pragma unroll
        for (int rs = 0; rs < 2; ++rs) {
            const double v = plane[(size_t)(g.ny - 1 + rs) * g.ld + c];
            inbox_row(a.inbox[tr] + so, a.ib, fo + rs)[ct] = v;
            if (ct == 1) {
                const int t2 = (tr + g.nranks - 1) % g.nranks;
                inbox_row(a.inbox[t2] + so, a.ib, fo + rs)[g.start[t2 + 1] - g.start[t2] + 1] = v;
            }
            if (ct == wt) inbox_row(a.inbox[(tr + 1) % g.nranks] + so, a.ib, fo + rs)[0] = v;
        }
    }
}

// center fields only (loc 0): ghost row = sgn * mirrored top row, E-W ghosts, fill on closed edges
struct NegRanges { int lo[2], hi[2]; };
__global__ void halo_apply_multi_kernel(double *plane0, size_t plane_stride, Geom g, InboxGeom ib, double *inbox0,
                                        int first, NegRanges neg, double fill, int west_closed, int east_closed)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
    const int pa = first + p;                                      // plane number in the whole stack
    const double sgn = ((pa >= neg.lo[0] && pa < neg.hi[0]) || (pa >= neg.lo[1] && pa < neg.hi[1])) ? -1.0 : 1.0;
    double *plane = plane0 + (size_t)p * plane_stride;
    double *inbox = inbox0 + (size_t)(p >> 1) * inbox_doubles(ib);
    const int fo = (p & 1) * 2;
    const int nxb = g.nxl + 2;
    if (t < nxb) {
        const int c = t;
        plane[c] = fill;                                           // south ghost row
        plane[(size_t)(g.ny + 1) * g.ld + c] = g.tripole ? sgn * inbox_row(inbox, ib, fo + 1)[c] : fill;
    } else if (t < nxb + g.ny) {
        const int r = t - nxb + 1;
        plane[(size_t)r * g.ld] = west_closed ? fill : inbox_col(inbox, ib, fo)[r];
        plane[(size_t)r * g.ld + g.nxl + 1] = east_closed ? fill : inbox_col(inbox, ib, fo + 1)[r];
    }
}

// ice_HaloUpdate_stress: ghost row of a1 <- mirrored top physical row of a2 (center, +1);
// the closed south row (and closed E-W ghosts) get 0 (:3351-3353).
__global__ void halo_stress_apply_kernel(double *a1, Geom g, InboxGeom ib, double *inbox,
                                         int west_closed, int east_closed)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int nxb = g.nxl + 2;
    if (t < nxb) {
        a1[t] = 0.0;
        if (g.tripole) a1[(size_t)(g.ny + 1) * g.ld + t] = inbox_row(inbox, ib, 1)[t];
    } else if (t < nxb + g.ny + 2) {
        const int r = t - nxb;
        if (west_closed) a1[(size_t)r * g.ld] = 0.0;
        if (east_closed) a1[(size_t)r * g.ld + g.nxl + 1] = 0.0;
    }
}

// ---- host <-> device staging: the caller's arrays are dense (nx_block x ny_block), the arena's planes are
// pitched (rows on 128-byte lines for TMA).  One contiguous copy per array over the host link, then ONE kernel
// that moves all arrays of the batch between the dense staging area and their planes (round 1 issued a
// row-pitched 2-D copy per array: 3.6 kB rows at N = 8 ran at 13 GB/s).
constexpr int kRepackMax = 48;
struct RepackArgs {
    int n, nxb, nyb, ld;
    void *pitched[kRepackMax];          // plane in the arena (double or int32)
    void *dense[kRepackMax];            // staging area
    unsigned char is_int[kRepackMax];
};
template <bool TO_PLANES>
__global__ void repack_kernel(const __grid_constant__ RepackArgs a)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y, k = blockIdx.z;
    if (c >= a.nxb) return;
    const size_t ip = (size_t)r * a.ld + c, id = (size_t)r * a.nxb + c;
    if (a.is_int[k]) {
        int *p = static_cast<int *>(a.pitched[k]), *d = static_cast<int *>(a.dense[k]);
        if (TO_PLANES) p[ip] = d[id]; else d[id] = p[ip];
    } else {
        double *p = static_cast<double *>(a.pitched[k]), *d = static_cast<double *>(a.dense[k]);
        if (TO_PLANES) p[ip] = d[id]; else d[id] = p[ip];
    }
}

// restart hand-off (source/ice_restart_driver.F90:431-466): the unused south ghost row and west ghost column
// of the 12 stresses are zeroed after a restart file has been read
__global__ void zero_sw_ghosts_kernel(double *s0, size_t plane_elems, int nxb, int nyb, int ld)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x, q = blockIdx.y;
    double *s = s0 + (size_t)q * plane_elems;
    if (t < nxb) s[t] = 0.0;                                   // j = 1 (nghost = 1)
    else if (t < nxb + nyb) s[(size_t)(t - nxb) * ld] = 0.0;   // i = 1
}

// ---- cross-rank synchronisation: SyncArgs, peer_signal, peer_wait live in evp_kernels.cuh ---------
__global__ void halo_signal_kernel(const __grid_constant__ SyncArgs a)
{
    if (threadIdx.x < 32) peer_signal_warp(a, threadIdx.x);
}

__global__ void halo_wait_kernel(const __grid_constant__ SyncArgs a)
{
    peer_wait(a, threadIdx.x);
    __syncthreads();
    __threadfence_system();
}

// ---- evp_prep1 -------------------------------------------------------------------------------
struct Prep1Args {
    Geom g;
    double rhoi, rhos, a_min, m_min;
    const double *aice, *vice, *vsno, *strairxT, *strairyT;
    const int *tmask;
    double *tmass, *work1, *work2, *icetmask;   // work1/2 <- strairxT/yT (for t2ugrid_vector)
};

__device__ __forceinline__ bool tmphm_at(const Prep1Args &a, size_t q)
{
    if (!a.tmask[q]) return false;
    const double tm = a.rhoi * a.vice[q] + a.rhos * a.vsno[q];
    return (a.aice[q] > a.a_min) && (tm > a.m_min);
}

__global__ void prep1_kernel(const __grid_constant__ Prep1Args a)
{
    const Geom &g = a.g;
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c > g.nxl + 1) return;
    const size_t q = (size_t)r * g.ld + c;
    a.tmass[q] = a.tmask[q] ? (a.rhoi * a.vice[q] + a.rhos * a.vsno[q]) : 0.0;
    a.work1[q] = a.strairxT[q];
    a.work2[q] = a.strairyT[q];
    double m = 0.0;
    if (c >= 1 && c <= g.nxl && r >= 1 && r <= g.ny) {
        bool any = false;
#pragma unroll
        for (int dr = -1; dr <= 1; ++dr)
#pragma unroll
            for (int dc = -1; dc <= 1; ++dc) any |= tmphm_at(a, q + (ptrdiff_t)dr * g.ld + dc);
        if (any && a.tmask[q]) m = 1.0;
    }
    a.icetmask[q] = m;
}

// ---- to_ugrid x4 -----------------------------------------------------------------------------
struct ToUArgs {
    Geom g;
    const double *tarea, *uarea;
    const double *in[4];
    double *out[4];
};

__global__ void to_ugrid4_kernel(const __grid_constant__ ToUArgs a)
{
    const Geom &g = a.g;
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c > g.nxl + 1) return;
    const size_t q = (size_t)r * g.ld + c;
    const bool phys = c >= 1 && c <= g.nxl && r >= 1 && r <= g.ny;
#pragma unroll
    for (int f = 0; f < 4; ++f) {
        double w = 0.0;
        if (phys) {
            const double *x = a.in[f];
            w = 0.25 * (x[q] * a.tarea[q] + x[q + 1] * a.tarea[q + 1] + x[q + g.ld] * a.tarea[q + g.ld] +
                        x[q + g.ld + 1] * a.tarea[q + g.ld + 1]) / a.uarea[q];
        }
        a.out[f][q] = w;
    }
}

// ---- evp_prep2 (+ ice_strength masking, + the loop's active-cell mask) --------------------------
struct Prep2Args {
    Geom g;
    double dt, a_min, m_min, cosw, sinw, gravit, rhow;
    int tilt_mode, strength_masked;
    int revised;                                // revised EVP: all stresses restart from zero (:492-504)
    const double *Cw;
    double *vrelc;                              // aiu*rhow*Cw: the constant factor of stepu's vrel
    const double *aiu, *umass, *fcor, *uocn, *vocn, *strairx, *strairy, *ss_tltx, *ss_tlty;
    const double *icetmask, *strength_in;
    const int *umask;
    int *iceumask;
    double *fm, *strtltx, *strtlty, *strocnx, *strocny, *strintx, *strinty;
    double *umassdti, *waterx, *watery, *forcex, *forcey, *strength;
    double *sA[12], *sB[12];                    // both ping-pong copies of the stresses
    double *uA, *vA, *uvel_init, *vvel_init;
    unsigned char *mask;
    int ldm;
    unsigned long long *counts;                 // [0] active T cells, [1] active U cells
};

__global__ void prep2_kernel(const __grid_constant__ Prep2Args a)
{
    const Geom &g = a.g;
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c > g.nxl + 1) return;
    const size_t q = (size_t)r * g.ld + c;
    const bool tice = a.icetmask[q] == 1.0;
    double waterx = 0.0, watery = 0.0, forcex = 0.0, forcey = 0.0, umassdti = 0.0;   // :486-490
    if (!tice || a.revised) {                                                          // :492-517
#pragma unroll
        for (int s = 0; s < 12; ++s) { a.sA[s][q] = 0.0; a.sB[s][q] = 0.0; }
    } else {
#pragma unroll
        for (int s = 0; s < 12; ++s) a.sB[s][q] = a.sA[s][q];   // cells the loop never stores
    }
    unsigned m = 0;
    if (tice && c >= 1 && r >= 1) m |= 1u;                                           // :528-537
    const bool phys = c >= 1 && c <= g.nxl && r >= 1 && r <= g.ny;
    double st = 0.0;
    if (phys) {
        const bool old = a.iceumask[q] != 0;
        const bool ice = a.umask[q] && (a.aiu[q] > a.a_min) && (a.umass[q] > a.m_min);  // :551
        a.iceumask[q] = ice ? 1 : 0;
        double u = a.uA[q], v = a.vA[q];
        if (ice) {
            m |= 2u;
            if (!old) { u = a.uocn[q]; v = a.vocn[q]; }                                // :560-563
            umassdti = a.umass[q] / a.dt;                                              // :587
            const double fm = a.fcor[q] * a.umass[q];
            const double sg = copysign(1.0, fm);
            a.fm[q] = fm;
            waterx = a.uocn[q] * a.cosw - a.vocn[q] * a.sinw * sg;
            watery = a.vocn[q] * a.cosw + a.uocn[q] * a.sinw * sg;
            double tx, ty;
            if (a.tilt_mode == 0) { tx = -fm * a.vocn[q]; ty = fm * a.uocn[q]; }
            else { tx = -a.gravit * a.umass[q] * a.ss_tltx[q]; ty = -a.gravit * a.umass[q] * a.ss_tlty[q]; }
            a.strtltx[q] = tx; a.strtlty[q] = ty;
            forcex = a.strairx[q] + tx;
            forcey = a.strairy[q] + ty;
        } else {
            u = 0.0; v = 0.0;                                                          // :566-571
            a.strintx[q] = 0.0; a.strinty[q] = 0.0; a.strocnx[q] = 0.0; a.strocny[q] = 0.0;
        }
        a.uA[q] = u; a.vA[q] = v;
        a.uvel_init[q] = u; a.vvel_init[q] = v;                                       // :574-575
        st = (!a.strength_masked || tice) ? a.strength_in[q] : 0.0;
    }
    a.strength[q] = st;
    a.waterx[q] = waterx; a.watery[q] = watery; a.forcex[q] = forcex; a.forcey[q] = forcey;
    a.umassdti[q] = umassdti;
    // stepu evaluates vrel = aiu*rhow*Cw*sqrt(...) left to right (ice_dyn_shared.F90:708-709): the
    // first two products do not change inside the loop -> one plane instead of two
    a.vrelc[q] = a.aiu[q] * a.rhow * a.Cw[q];
    a.mask[(size_t)r * a.ldm + c] = (unsigned char)m;
    // active-cell counts (metric denominator); one atomic per warp
    const unsigned bt = __ballot_sync(__activemask(), m & 1u), bu = __ballot_sync(__activemask(), m & 2u);
    if ((threadIdx.x & 31) == 0) {
        if (bt) atomicAdd(&a.counts[0], (unsigned long long)__popc(bt));
        if (bu) atomicAdd(&a.counts[1], (unsigned long long)__popc(bu));
    }
}

// the T mask bit of the ghost cells depends on icetmask AFTER its halo update, which prep2 sees;
// but strength's ghost values come from its own halo: nothing else to do here.

// ---- evp_finish ------------------------------------------------------------------------------
struct FinishArgs {
    Geom g;
    double rhow, cosw, sinw;
    const double *u, *v, *uocn, *vocn, *aiu, *fm, *Cw;
    const int *iceumask;
    double *strocnx, *strocny, *work1, *work2;      // work1/2 <- strocnxT/yT on the U grid
};

__global__ void finish_kernel(const __grid_constant__ FinishArgs a)
{
    const Geom &g = a.g;
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c > g.nxl + 1) return;
    const size_t q = (size_t)r * g.ld + c;
    double tx = 0.0, ty = 0.0;                                                         // :806-811
    if (c >= 1 && c <= g.nxl && r >= 1 && r <= g.ny && a.iceumask[q]) {
        const double du = a.uocn[q] - a.u[q], dv = a.vocn[q] - a.v[q];
        double vrel = a.rhow * a.Cw[q] * sqrt(du * du + dv * dv);                      // :818
        vrel = vrel * a.aiu[q];                                                        // :827
        const double sg = copysign(1.0, a.fm[q]);
        const double sx = vrel * (du * a.cosw - dv * a.sinw * sg);
        const double sy = vrel * (dv * a.cosw + du * a.sinw * sg);
        a.strocnx[q] = sx; a.strocny[q] = sy;
        tx = sx / a.aiu[q]; ty = sy / a.aiu[q];                                        // :840-841
    }
    a.work1[q] = tx; a.work2[q] = ty;
}

// to_tgrid x2: out = 0 outside the physical cells (evp_finish zero-fill, :806-811)
__global__ void to_tgrid2_kernel(Geom g, const double *tarea, const double *uarea, const double *in0,
                                 const double *in1, double *out0, double *out1)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c > g.nxl + 1) return;
    const size_t q = (size_t)r * g.ld + c;
    double w0 = 0.0, w1 = 0.0;
    if (c >= 1 && c <= g.nxl && r >= 1 && r <= g.ny) {
        const size_t w = q - 1, s = q - g.ld, sw = q - g.ld - 1;
        w0 = 0.25 * (in0[q] * uarea[q] + in0[w] * uarea[w] + in0[s] * uarea[s] + in0[sw] * uarea[sw]) / tarea[q];
        w1 = 0.25 * (in1[q] * uarea[q] + in1[w] * uarea[w] + in1[s] * uarea[s] + in1[sw] * uarea[sw]) / tarea[q];
    }
    out0[q] = w0; out1[q] = w1;
}

__global__ void double_to_int_kernel(const double *in, int *out, size_t n)
{
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q < n) out[q] = (int)in[q];
}

// ---- principal_stress (diagnostic for history; source/ice_dyn_shared.F90:853-893) -------------
// sig1, sig2 = (stressp_1 +- sqrt(stressm_1^2 + 4 stress12_1^2)) / 2 / prs_sig over the whole block,
// spval_dbl (1e30) where prs_sig <= puny.
__global__ void principal_stress_kernel(int nxb, int nyb, int ld, const double *stressp_1, const double *stressm_1,
                                        const double *stress12_1, const double *prs_sig, double *sig1, double *sig2)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c >= nxb || r >= nyb) return;
    const size_t q = (size_t)r * ld + c;
    const double p = prs_sig[q];
    double s1 = 1.0e30, s2 = 1.0e30;
    if (p > 1.0e-11) {
        const double root = sqrt(stressm_1[q] * stressm_1[q] + 4.0 * (stress12_1[q] * stress12_1[q]));
        s1 = (0.5 * (stressp_1[q] + root)) / p;
        s2 = (0.5 * (stressp_1[q] - root)) / p;
    }
    sig1[q] = s1;
    sig2[q] = s2;
}

}  // namespace evp
