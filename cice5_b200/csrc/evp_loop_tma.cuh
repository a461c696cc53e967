// evp_loop_tma.cuh -- the whole ndte-subcycle loop of evp(dt) as ONE persistent kernel (product path).
//
// reference: the ksub loop source/ice_dyn_evp.F90:336-410 -- stress (:520-849), stepu
// (source/ice_dyn_shared.F90:623-748) and the width-1 halo update of (uvel, vvel) (:392-400: E-W neighbours
// and the tripole u-fold, serial/ice_boundary.F90:809-890) -- for all ndte iterations in one launch.
//
// Same tiles, same TMA ring, same arithmetic as evp_subcycle_tma_kernel (evp_subcycle_tma.cuh): a work item
// is (strip of 124 U columns) x (chunk of rows) of ONE subcycle.  What is new is what separates two
// subcycles.  A launch per subcycle (round 1) pays, per subcycle, two launches, the queue prologue, a
// pipeline fill per CTA and a ragged end where every CTA waits for the slowest -- 95 us where the data
// motion needs 69 on the cells one GPU owns at N = 8 (profiles/r02_call1.md).  Here the only ordering is
// the one the data demands:
//
//   * positions of one global queue run over (subcycle, item) in subcycle-major order, items south to
//     north; persistent CTAs draw positions with one atomicAdd each;
//   * tile t may start subcycle k once the 3 x 3 tiles around it have finished subcycle k-1 (their new
//     velocities are what it reads, and they have finished reading the planes it is about to overwrite:
//     the ping-pong copy of subcycle k-2).  Every tile publishes a monotonic counter (st.release), the
//     producer warp of the CTA that drew the item polls the nine counters (ld.acquire) before its first
//     TMA load.  Because the queue is ordered, a tile's neighbours of the previous subcycle were drawn a
//     whole subcycle earlier: on grids with more tiles than CTAs nobody ever waits, the end of subcycle k
//     overlaps the start of k+1, and on small grids the kernel degenerates into a dataflow pipeline whose
//     period is one tile + one flag round trip;
//   * across GPUs NO fence at all: an edge strip sends its edge column to the neighbour rank as flag-in-data
//     words (ghost mailbox, below: every 64-bit store carries 32 bits of a velocity and the subcycle number, the
//     way NCCL's LL protocol does), and the neighbour's edge tile of the next subcycle polls the words in its OWN
//     memory and writes the values into its ghost column before its first TMA load.  (First version: ghost values
//     stored into the peer's planes + a system-scope release of a counter -- `st.release.sys` cost 20 us per edge
//     item under load, profiles/r02_scaling.md.)  Ranks may drift by almost a whole subcycle;
//   * the tripole fold of subcycle k is one more item at the end of the subcycle's positions: it waits
//     until the top tiles of every rank have finished k (one counter per rank, published by the CTA that
//     completes the count), folds rows ny, ny+1 of the new velocities and publishes `fold_flag`; only the
//     top tiles of subcycle k+1 -- drawn a whole subcycle later -- depend on it.
//
// No CTA ever waits for a CTA that has not started (a position only depends on earlier positions), so the
// kernel needs no co-residency guarantee and no cooperative launch.  It must not share a device with a
// peer rank's kernel queued behind it in the same stream (evp_gpu_compute_group keeps the launch-per-
// subcycle path for that reason).
#pragma once
#include "evp_subcycle_tma.cuh"

namespace evp {

constexpr int kItemFold = 0x20000000;        // tag of the fold item of a subcycle
struct LoopTmaArgs {
    LoopArgs a;                  // geometry, constants, last-subcycle outputs, uvel_init (state pointers unused)
    int plane_state[2];          // first plane of ping-pong copy A / B
    int plane_t, plane_u;
    double *state[2];            // plane(S_UA), plane(S_UB): u, v, 12 stresses, plane_elems apart
    double *west_state[2], *east_state[2];   // the E-W neighbours' (peer memory; self when cyclic on one rank; null: closed)
    size_t plane_elems;
    double *inbox[2][kMaxRanks]; // tripole inboxes of every rank, both buffers
    int parity0;                 // copy that holds the state before subcycle 0
    int inbox0;                  // inbox buffer of subcycle 0 (alternates)
    int strips, chunks;
    int nsteps;                  // subcycles of this launch (ndte)
    int has_fold;                // tripole: one fold item per subcycle
    int *queue;                  // one counter for the whole launch
    const int *item_list, *item_count;       // [0] items with ice, [2] of them top tiles (rend >= ny)
    const unsigned char *tile_flags;
    int flags_ld;
    int water_is_ocn;
    double puny;
    // dependency counters (u64, subcycles completed): mine [chunk * strips + strip]; the west / east
    // neighbour's edge strip [chunk] (written by the neighbour; own strips when cyclic on one rank)
    unsigned long long *tflag, *wflag, *eflag;
    int ew_self;                             // cyclic on one rank: the E-W neighbours are my own edge strips
    unsigned long long *pub_w, *pub_e;       // where my edge strips publish: the neighbours' eflag / wflag (null: none)
    unsigned long long *topflag;             // [rank]: subcycles whose top tiles are complete on that rank (mine to read)
    unsigned long long *pub_top[kMaxRanks];  // every rank's topflag array
    unsigned long long *fold_flag;           // subcycles folded on this rank
    unsigned long long *top_count;           // top tiles finished on this rank (all subcycles)
    int rank, nranks;
    int *error;
    unsigned long long timeout_ns;
    // ghost mailbox (nranks > 1): [kLLBufs][side: 0 = my west ghost, 1 = my east ghost][row 0 .. nyb-1][4 words]
    unsigned long long *ll_my;               // mine: the neighbours write it, I poll it
    unsigned long long *ll_w, *ll_e;         // the west / east neighbour's (peer memory); null: no neighbour on that side
    int ll_on, nyb;
};

// ---- ghost mailbox: flag-in-data exchange of the E-W ghost columns between ranks -------------------------
// One row of one side = 4 words {u low, u high, v low, v high}; word = (32 bits of payload << 32) | flag, flag =
// subcycle + 1 (never 0; the mailbox is zeroed at the start of a call), bit 31 = "row not computed this call: keep the
// value you have".  A naturally aligned 64-bit store is single-copy atomic, so a word whose flag matches carries
// its payload: no fence, no separate counter, and both sides only use strong (volatile) accesses -- no data race in
// the PTX model.  Buffer = subcycle & 3: a slot is rewritten 4 subcycles later, by when the receiver has long
// consumed it (its reader must have finished before the writer's own dependencies two subcycles on can be met).
constexpr int kLLBufs = 4;
constexpr unsigned kLLUnchanged = 0x80000000u;

__device__ __forceinline__ volatile unsigned long long *ll_row(unsigned long long *box, int nyb, int step, int side, int row)
{
    return box + ((((size_t)(step & (kLLBufs - 1))) * 2 + side) * nyb + row) * 4;
}
__device__ __forceinline__ void ll_push(unsigned long long *box, int nyb, int step, int side, int row, double u, double v,
                                        bool computed)
{
    volatile unsigned long long *w = ll_row(box, nyb, step, side, row);
    const unsigned long long flag = (unsigned long long)(unsigned)(step + 1) | (computed ? 0u : kLLUnchanged);
    const unsigned long long ub = (unsigned long long)__double_as_longlong(u), vb = (unsigned long long)__double_as_longlong(v);
    w[0] = (ub << 32) | flag;
    w[1] = (ub & 0xffffffff00000000ull) | flag;
    w[2] = (vb << 32) | flag;
    w[3] = (vb & 0xffffffff00000000ull) | flag;
}
// wait for the four words of (step, side, row); false: the sender did not compute the row (or the wait gave up)
__device__ __forceinline__ bool ll_take(unsigned long long *box, int nyb, int step, int side, int row, double &u, double &v,
                                        int *error, unsigned long long timeout_ns)
{
    volatile unsigned long long *w = ll_row(box, nyb, step, side, row);
    const unsigned want = (unsigned)(step + 1);
    unsigned long long x[4];
    unsigned long long t0 = 0;
    for (int q = 0; q < 4; ++q) {
        for (;;) {
            x[q] = w[q];
            if (((unsigned)x[q] & ~kLLUnchanged) == want) break;
            if (t0 == 0) t0 = global_ns();
            __nanosleep(32);
            if (*(volatile int *)error) return false;
            if (global_ns() - t0 > timeout_ns) { *(volatile int *)error = 1; return false; }
        }
    }
    if ((unsigned)x[0] & kLLUnchanged) return false;
    u = __longlong_as_double((long long)((x[0] >> 32) | (x[1] & 0xffffffff00000000ull)));
    v = __longlong_as_double((long long)((x[2] >> 32) | (x[3] & 0xffffffff00000000ull)));
    return true;
}

// After the loop kernel (and the cross-rank barrier behind it): the ghost columns of the final planes from the
// mailbox rows of the last subcycle of the loop -- inside the loop a receiver tile does this for itself one subcycle
// later.  One thread per (row, side); rows 1 .. jhi.
__global__ void ll_drain_kernel(double *u, size_t plane_elems, int ld, int nxl, int nyb, unsigned long long *box, int step,
                                int jhi, int recv_w, int recv_e)
{
    const int j = 1 + blockIdx.x * blockDim.x + threadIdx.x, side = blockIdx.y;
    if (j > jhi || !(side == 0 ? recv_w : recv_e)) return;
    volatile unsigned long long *w = ll_row(box, nyb, step, side, j);
    unsigned long long x[4];
    for (int q = 0; q < 4; ++q) x[q] = w[q];
    for (int q = 0; q < 4; ++q)
        if ((unsigned)x[q] != (unsigned)(step + 1)) return;           // not sent, or "keep the value you have"
    const size_t at = (size_t)j * ld + (side == 0 ? 0 : nxl + 1);
    u[at] = __longlong_as_double((long long)((x[0] >> 32) | (x[1] & 0xffffffff00000000ull)));
    u[plane_elems + at] = __longlong_as_double((long long)((x[2] >> 32) | (x[3] & 0xffffffff00000000ull)));
}

// SYS: the counter is written or read by another GPU (peer memory); otherwise gpu scope is enough and cheaper
template <bool SYS>
__device__ __forceinline__ void st_release_u64(unsigned long long *p, unsigned long long v)
{
    if (SYS) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    else asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
template <bool SYS>
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p)
{
    unsigned long long v;
    if (SYS) asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    else asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// spin until *p >= want; gives up when another waiter has timed out or after timeout_ns (error word set:
// the host returns EVP_ERR_PEER, the results of the call are void)
template <bool SYS>
__device__ __forceinline__ void wait_counter(const unsigned long long *p, unsigned long long want, int *error,
                                             unsigned long long timeout_ns)
{
    if (ld_acquire_u64<SYS>(p) >= want) return;
    const unsigned long long t0 = global_ns();
    for (;;) {
        __nanosleep(32);
        if (ld_acquire_u64<SYS>(p) >= want) return;
        if (*(volatile int *)error) return;
        if (global_ns() - t0 > timeout_ns) { *(volatile int *)error = 1; return; }
    }
}

// counter of the tile (strip + ds, chunk + dc) as seen from this rank, or null when there is none
__device__ __forceinline__ const unsigned long long *neighbour_counter(const LoopTmaArgs &p, int strip, int chunk,
                                                                       bool &remote)
{
    remote = false;
    if (chunk < 0 || chunk >= p.chunks) return nullptr;
    if (strip < 0) {
        if (p.ew_self) return p.tflag + (size_t)chunk * p.strips + (p.strips - 1);
        remote = true;
        return (p.wflag && !p.ll_on) ? p.wflag + chunk : nullptr;
    }
    if (strip >= p.strips) {
        if (p.ew_self) return p.tflag + (size_t)chunk * p.strips;
        remote = true;
        return (p.eflag && !p.ll_on) ? p.eflag + chunk : nullptr;
    }
    return p.tflag + (size_t)chunk * p.strips + strip;
}

// The row sweep of one work item (strip, rows r0 .. rend) of one subcycle by the consumer warps: stress on T rows
// r0 .. rend, stepu on U rows r0 .. rend-1, new state into copy par^1 (+ the neighbours' ghost columns and the
// tripole inboxes).  Same statements as the consumer loop of evp_subcycle_tma_kernel.
template <bool LAST, bool REVISED>
__device__ __forceinline__ void sweep_item(const LoopTmaArgs &p, unsigned char *ring, unsigned long long *full_bar,
                                           unsigned long long *empty_bar, int &stage, unsigned &phase, int strip, int r0,
                                           int rend, int par, int ibuf, int step, int warp, int lane)
{
    const LoopArgs &a = p.a;
    const Consts &k = a.k;
    const int ld = a.ld;
    const size_t pe = p.plane_elems;
    constexpr int kMetaOff = (kStateDepth + kTDepth + kUDepth) * kTmaBoxW * 8;
    const int bi = 1 + warp * kUColsPerWarp + lane;                     // my column inside the box
    double *const snew = p.state[par ^ 1];
    double *const u1 = snew, *const v1 = snew + pe;
    const int c = 1 + strip * kTmaUCols + warp * kUColsPerWarp + lane;   // T column (U column if lane < 31)
    const bool warp_on = (c - lane) <= a.nxl;                       // warp-uniform
    const int mshift = (strip * kTmaUCols) & 15;                    // mask box starts this many bytes early
    const bool tcol = c <= a.nxl + 1;
    const bool ucol = lane < kUColsPerWarp && c <= a.nxl;
    const bool tstore_col = tcol && (lane < kUColsPerWarp || c == a.nxl + 1);
    double *const uW = p.west_state[par ^ 1], *const uE = p.east_state[par ^ 1];

    // row r0-1: velocities only
    double u_s, v_s, uw_s, vw_s;
    double htn_s = 0.0;                                             // HTN of the row below (compact rows)
    bool loaded_s;                                                  // state of the row below was fetched
    {
        const double *st = reinterpret_cast<const double *>(ring + (size_t)stage * kStageBytes);
        const unsigned meta = *reinterpret_cast<const unsigned *>(
            ring + (size_t)stage * kStageBytes + kMetaOff + kMaskFlagOff);
        loaded_s = (meta & kMetaState) != 0u;
        if (meta & kMetaHtnExtra) htn_s = st[(kStateDepth + kSlotHtnExtra) * kTmaBoxW + bi];
        u_s = st[bi]; v_s = st[kTmaBoxW + bi]; uw_s = st[bi - 1]; vw_s = st[kTmaBoxW + bi - 1];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[stage]);
        if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
    }
    double s1p = 0.0, s5p = 0.0, n2p = 0.0, n7p = 0.0;
    unsigned mprev = 0;

    for (int r = r0; r <= rend; ++r) {
        mbar_wait(&full_bar[stage], phase);
        const double *st = reinterpret_cast<const double *>(ring + (size_t)stage * kStageBytes);
        const double *tc = st + kStateDepth * kTmaBoxW;
        const double *uc = tc + kTDepth * kTmaBoxW;
        const unsigned char *mk = reinterpret_cast<const unsigned char *>(uc + kUDepth * kTmaBoxW);
        const unsigned meta = *reinterpret_cast<const unsigned *>(mk + kMaskFlagOff);
        const bool loaded_c = (meta & kMetaState) != 0u;
        const double u_c = st[bi], v_c = st[kTmaBoxW + bi];
        const double uw_c = st[bi - 1], vw_c = st[kTmaBoxW + bi - 1];
        const int mpos = bi + mshift;                               // mask box b covers bytes [128 b, 128 b + 144)
        const int mbox = min(mpos >> 7, kMaskBoxes - 1);            // and sits 256 b bytes into the mask area
        const unsigned m = (warp_on && tcol) ? mk[mbox * 256 + (mpos - mbox * 128)] : 0u;
        const bool tact = (m & 1u) != 0;
        const bool uact = warp_on && ucol && (mprev & 2u) && r > r0;
        const bool any_t = __ballot_sync(0xffffffffu, tact) != 0u;
        double s[12], tco[11], uco[kUDepth];
        // tco: dxt dyt dxhy dyhx cxp cyp cxm cym tinyarea strength tarear
        if (any_t) {
#pragma unroll
            for (int q = 0; q < 12; ++q) s[q] = st[(2 + q) * kTmaBoxW + bi];
            if (meta & kMetaCompact) {
                // metric terms from HTE / HTN by the reference's own formulas (ice_grid.F90:354-367,1456,1536);
                // set_grid checked that they reproduce the host arrays bit for bit on this row segment
                const double hte = tc[bi], hte_w = tc[bi - 1], htn = tc[kTmaBoxW + bi];
                metric_terms(hte, hte_w, htn, htn_s, tco);
                tco[9] = tc[2 * kTmaBoxW + bi];
                tco[8] = (meta & kMetaTinyFromMetrics) ? __dmul_rn(p.puny, __dmul_rn(tco[0], tco[1]))
                                                       : tc[3 * kTmaBoxW + bi];
            } else {
                tco[9] = tc[bi]; tco[8] = tc[kTmaBoxW + bi];
#pragma unroll
                for (int q = 0; q < 8; ++q) tco[q] = tc[(2 + q) * kTmaBoxW + bi];
            }
            if (LAST) tco[10] = tc[kSlotTarear * kTmaBoxW + bi];
        }
        // southern HTN of the next row
        if ((meta & (kMetaCompact | kMetaT)) == (kMetaCompact | kMetaT)) htn_s = tc[kTmaBoxW + bi];
        else if (meta & kMetaHtnExtra) htn_s = tc[kSlotHtnExtra * kTmaBoxW + bi];

        const size_t idx = (size_t)r * ld + c;
        double str[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        if (tact) {
            double divu, shear, rdg_conv, rdg_shear, prs_sig;
            stress_cell<LAST>(k, u_c, uw_c, u_s, uw_s, v_c, vw_c, v_s, vw_s, tco[0], tco[1], tco[2], tco[3],
                              tco[4], tco[5], tco[6], tco[7], tco[8], tco[9], LAST ? tco[10] : 0.0, s, str, divu,
                              shear, rdg_conv, rdg_shear, prs_sig);
            if (tstore_col && (r < rend || r == a.ny + 1)) {
                double *sp = snew + 2 * pe + idx;
#pragma unroll
                for (int q = 0; q < 12; ++q) sp[q * pe] = s[q];
                if (LAST) {
                    a.divu[idx] = divu; a.shear[idx] = shear; a.rdg_conv[idx] = rdg_conv;
                    a.rdg_shear[idx] = rdg_shear; a.prs_sig[idx] = prs_sig;
                }
            }
        }
        const double e2 = shfl_down1(str[1]), e4 = shfl_down1(str[3]);
        const double e7 = shfl_down1(str[6]), e8 = shfl_down1(str[7]);

        if (r > r0 && warp_on && loaded_s) {                        // stepu on U row r-1 (:700-746)
            const size_t iu = idx - ld;
            double unew = u_s, vnew = v_s;
            if (uact) {
#pragma unroll
                for (int q = 0; q < kUDepth - 2; ++q) uco[q] = uc[q * kTmaBoxW + bi];
                uco[8] = p.water_is_ocn ? uco[1] : uc[8 * kTmaBoxW + bi];
                uco[9] = p.water_is_ocn ? uco[2] : uc[9 * kTmaBoxW + bi];
                // uco: aiu*rhow*Cw uocn vocn forcex forcey umassdti fm uarear waterx watery
                const double uold = u_s, vold = v_s;
                const double du = uco[1] - uold, dv = uco[2] - vold;
                const double vrel = uco[0] * sqrt(du * du + dv * dv);
                const double taux = vrel * uco[8], tauy = vrel * uco[9];
                const double umdt = uco[5], fm = uco[6];
                const double cca = (k.brlx + k.revp) * umdt + vrel * k.cosw;
                const double ccb = fm + copysign(1.0, fm) * vrel * k.sinw;
                const double ab2 = cca * cca + ccb * ccb;
                const double uar = uco[7];
                const double strintx = uar * (s1p + n2p + str[2] + e4);
                const double strinty = uar * (s5p + str[5] + n7p + e8);
                // revised EVP (revp = 1) also needs uvel_init (:731-734): two direct global loads,
                // only in the REVISED instantiation
                const double cc1 = strintx + uco[3] + taux +
                                   umdt * (REVISED ? k.brlx * uold + k.revp * a.uvel_init[iu] : k.brlx * uold);
                const double cc2 = strinty + uco[4] + tauy +
                                   umdt * (REVISED ? k.brlx * vold + k.revp * a.vvel_init[iu] : k.brlx * vold);
                unew = (cca * cc1 + ccb * cc2) / ab2;
                vnew = (cca * cc2 - ccb * cc1) / ab2;
                if (LAST) { a.strintx[iu] = strintx; a.strinty[iu] = strinty; }
            }
            if (ucol) {
                u1[iu] = unew; v1[iu] = vnew;
                if (p.ll_on) {                                      // my edge columns are the neighbours' ghosts
                    if (c == 1 && p.ll_w) ll_push(p.ll_w, p.nyb, step, 1, r - 1, unew, vnew, true);
                    if (c == a.nxl && p.ll_e) ll_push(p.ll_e, p.nyb, step, 0, r - 1, unew, vnew, true);
                } else {
                    if (c == 1 && uW) {
                        const size_t ig = (size_t)(r - 1) * ld + a.nxl_west + 1;
                        uW[ig] = unew; uW[pe + ig] = vnew;
                    }
                    if (c == a.nxl && uE) {
                        const size_t ig = (size_t)(r - 1) * ld;
                        uE[ig] = unew; uE[pe + ig] = vnew;
                    }
                }
                if (a.tripole && r - 1 >= a.ny - 1)
                    tripole_push(a, p.inbox[ibuf], c, (r - 1) - (a.ny - 1), unew, vnew);
            }
        } else if (r > r0 && p.ll_on && ucol) {
            // U row r-1 of this segment is not computed in this call (no ice): the neighbour must not wait for it
            if (c == 1 && p.ll_w) ll_push(p.ll_w, p.nyb, step, 1, r - 1, 0.0, 0.0, false);
            if (c == a.nxl && p.ll_e) ll_push(p.ll_e, p.nyb, step, 0, r - 1, 0.0, 0.0, false);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[stage]);              // row done: slot free
        if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
        s1p = str[0]; s5p = str[4]; n2p = e2; n7p = e7;
        u_s = u_c; v_s = v_c; uw_s = uw_c; vw_s = vw_c; mprev = m; loaded_s = loaded_c;
    }

}

template <bool REVISED>
__global__ void __launch_bounds__((kTmaConsumers + 1) * 32, kTmaBlocksPerSM)
evp_loop_tma_kernel(const __grid_constant__ CUtensorMap mapS, const __grid_constant__ CUtensorMap mapS2,
                    const __grid_constant__ CUtensorMap mapT1, const __grid_constant__ CUtensorMap mapT3,
                    const __grid_constant__ CUtensorMap mapT4, const __grid_constant__ CUtensorMap mapT10,
                    const __grid_constant__ CUtensorMap mapU, const __grid_constant__ CUtensorMap mapM,
                    const __grid_constant__ LoopTmaArgs p)
{
    extern __shared__ __align__(1024) unsigned char ring[];
    __shared__ unsigned long long full_bar[kTmaStages], empty_bar[kTmaStages];
    const LoopArgs &a = p.a;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int kMetaOff = (kStateDepth + kTDepth + kUDepth) * kTmaBoxW * 8;   // mask bytes + stage words

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kTmaStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], kTmaConsumers); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    int stage = 0;
    unsigned phase = 0;
    const int nkept = p.item_count[0];
    const int per_step = nkept + (p.has_fold ? 1 : 0);
    const long long total = (long long)per_step * p.nsteps;

    if (warp == kTmaConsumers) {
        // ------------------------------ producer warp ------------------------------
        // lane 0 always holds the NEXT position already (drawn while the previous item was being issued): the
        // round trip of the atomic is off the path between two items
        int qnext = 0;
        if (lane == 0) qnext = atomicAdd(p.queue, 1);
        for (;;) {
            const int q = __shfl_sync(0xffffffffu, qnext, 0);
            if (lane == 0 && q < total) qnext = atomicAdd(p.queue, 1);
            if (q >= total || per_step == 0) {                          // queue drained: tell the consumers
                if (lane == 0) {
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    *reinterpret_cast<volatile int *>(ring + (size_t)stage * kStageBytes + kMetaOff + kMaskFlagOff + 4) = -1;
                    mbar_arrive(&full_bar[stage]);
                }
                break;
            }
            const int k = q / per_step;                                 // subcycle, 0-based
            const int pos = q - k * per_step;
            if (pos == nkept) {                                         // the fold of subcycle k
                if (lane == 0) {
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    unsigned char *mk = ring + (size_t)stage * kStageBytes + kMetaOff;
                    *reinterpret_cast<volatile int *>(mk + kMaskFlagOff + 4) = kItemFold;
                    *reinterpret_cast<volatile int *>(mk + kMaskFlagOff + 8) = k;
                    mbar_arrive(&full_bar[stage]);
                    if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
                }
                continue;
            }
            const int item = p.item_list[pos];
            const int strip = item % p.strips, chunk = item / p.strips;
            int r0, rend;
            chunk_rows(chunk, a.rows_per_chunk, a.ny, r0, rend);
            if (k > 0) {
                // dependencies: the 3 x 3 tiles around this one have finished subcycle k-1 (lanes 0..8 poll
                // one counter each), and, for a tile that reads rows ny / ny+1, the fold of k-1 (lane 9)
                if (lane < 9) {
                    bool remote;
                    const unsigned long long *f = neighbour_counter(p, strip + lane % 3 - 1, chunk + lane / 3 - 1, remote);
                    if (f) {
                        if (remote) wait_counter<true>(f, (unsigned long long)k, p.error, p.timeout_ns);
                        else wait_counter<false>(f, (unsigned long long)k, p.error, p.timeout_ns);
                    }
                } else if (lane == 9 && p.has_fold && rend >= a.ny) {
                    wait_counter<false>(p.fold_flag, (unsigned long long)k, p.error, p.timeout_ns);
                }
                if (p.ll_on) {
                    // ghost rows r0-1 .. rend of the OLD planes of this subcycle = what the neighbour's edge strip
                    // computed in subcycle k-1: out of my mailbox into my ghost column (strong stores: the tiles above
                    // and below write the rows we share with the same values).  Row ny of a tripole grid belongs to
                    // the fold item, row ny+1 to nobody.
                    double *const sold = p.state[p.parity0 ^ (k & 1)];
                    const int jlo = r0 - 1 > 1 ? r0 - 1 : 1, jtop = p.has_fold ? a.ny - 1 : a.ny;
                    const int jhi = rend < jtop ? rend : jtop;
                    for (int side = 0; side < 2; ++side) {
                        const unsigned long long *nbr = side == 0 ? p.wflag : p.eflag;
                        if (!nbr || (side == 0 ? strip != 0 : strip != p.strips - 1)) continue;
                        for (int j = jlo + lane; j <= jhi; j += 32) {
                            if (nbr[(j - 1) / a.rows_per_chunk] == ~0ull) continue;      // that tile never runs
                            double u, v;
                            if (ll_take(p.ll_my, p.nyb, k - 1, side, j, u, v, p.error, p.timeout_ns)) {
                                const size_t at = (size_t)j * a.ld + (side == 0 ? 0 : a.nxl + 1);
                                *(volatile double *)(sold + at) = u;
                                *(volatile double *)(sold + p.plane_elems + at) = v;
                            }
                        }
                    }
                }
                __syncwarp();
                // the planes were written through the generic proxy (other SMs, other GPUs); TMA reads them
                // through the async proxy
                asm volatile("fence.proxy.async.global;" ::: "memory");
            }
            if (lane == 0) {
                const int par = p.parity0 ^ (k & 1);
                const int plane_state = p.plane_state[par];
                const int x0 = strip * kTmaUCols;                       // box column 0 = U column c0 - 1
                const unsigned char *fl = p.tile_flags + (size_t)strip * p.flags_ld;
                for (int j = r0 - 1; j <= rend; ++j) {
                    // what this row segment needs (as in evp_subcycle_tma_kernel): T data if it has an active T
                    // cell; U data of the row below if that row has an active U cell; the state if this row or
                    // the next one has an active T cell or this row an active U cell.  The top rows are never
                    // skipped: their velocities feed the tripole inboxes every subcycle.
                    const bool top = j >= a.ny - 2;
                    const unsigned f0 = fl[j], f1 = fl[j + 1], fm1 = j > 0 ? fl[j - 1] : 0u;
                    const bool needT = j >= r0 && ((f0 & 1u) || top);
                    const bool needU = j > r0 && ((fm1 & 2u) || top);
                    const bool needS = ((f0 | f1) & 1u) || (f0 & 2u) || top;
                    const bool cmp = (f0 & 4u) != 0u, tinym = (f0 & 8u) != 0u;
                    const bool nextT = j + 1 <= rend && ((f1 & 1u) || j + 1 >= a.ny - 2) && (f1 & 4u);
                    const bool htn_extra = nextT && !(needT && cmp);
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    unsigned char *sb = ring + (size_t)stage * kStageBytes;
                    unsigned char *tb = sb + kStateDepth * kTmaBoxW * 8;
                    unsigned char *mk = sb + kMetaOff;
                    *reinterpret_cast<volatile unsigned *>(mk + kMaskFlagOff) =
                        (needS ? kMetaState : 0u) | (cmp ? kMetaCompact : 0u) | (tinym ? kMetaTinyFromMetrics : 0u) |
                        (htn_extra ? kMetaHtnExtra : 0u) | (needT ? kMetaT : 0u);
                    *reinterpret_cast<volatile int *>(mk + kMaskFlagOff + 4) = item;
                    *reinterpret_cast<volatile int *>(mk + kMaskFlagOff + 8) = k;
                    const int tdepth = cmp ? (tinym ? 3 : 4) : 10;
                    unsigned bytes = kMaskBoxW * kMaskBoxes;
                    const bool uv_only = j == r0 - 1;                    // first row of an item: velocities only
                    if (needS) bytes += (uv_only ? 2 : kStateDepth) * kTmaBoxW * 8;
                    if (needT) bytes += tdepth * kTmaBoxW * 8;
                    if (htn_extra) bytes += kTmaBoxW * 8;
                    if (needU) bytes += (p.water_is_ocn ? kUDepth - 2 : kUDepth) * kTmaBoxW * 8;
                    mbar_expect_tx(&full_bar[stage], bytes);
#pragma unroll
                    for (int mb = 0; mb < kMaskBoxes; ++mb)
                        tma_load_2d(mk + mb * 256, &mapM, &full_bar[stage], (x0 & ~15) + mb * 128, j);
                    if (needS) tma_load_3d(sb, uv_only ? &mapS2 : &mapS, &full_bar[stage], x0, j, plane_state);
                    if (needT) {
                        if (!cmp) tma_load_3d(tb, &mapT10, &full_bar[stage], x0, j, p.plane_t + 2);
                        else if (tinym) tma_load_3d(tb, &mapT3, &full_bar[stage], x0, j, p.plane_t);
                        else tma_load_3d(tb, &mapT4, &full_bar[stage], x0, j, p.plane_t);
                    }
                    if (htn_extra)
                        tma_load_3d(tb + kSlotHtnExtra * kTmaBoxW * 8, &mapT1, &full_bar[stage], x0, j, p.plane_t + 1);
                    if (needU)
                        tma_load_3d(sb + (kStateDepth + kTDepth) * kTmaBoxW * 8, &mapU, &full_bar[stage], x0, j - 1,
                                    p.plane_u);
                    if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
                }
            }
            __syncwarp();                                               // (only lane 0 tracks the ring position)
        }
        return;
    }

    // ------------------------------ consumer warps ------------------------------
    const Consts &k = a.k;
    const int ld = a.ld;
    const size_t pe = p.plane_elems;
    const int bi = 1 + warp * kUColsPerWarp + lane;                     // my column inside the box
    const int tid = warp * 32 + lane;
    for (;;) {
        mbar_wait(&full_bar[stage], phase);                             // first stage of the next item
        const unsigned char *mk0 = ring + (size_t)stage * kStageBytes + kMetaOff;
        const int item = *reinterpret_cast<const int *>(mk0 + kMaskFlagOff + 4);
        const int step = *reinterpret_cast<const int *>(mk0 + kMaskFlagOff + 8);
        if (item < 0) break;
        const int par = p.parity0 ^ (step & 1);
        double *const snew = p.state[par ^ 1];                          // u, v, 12 stresses of subcycle `step`
        double *const u1 = snew, *const v1 = snew + pe;
        const int ibuf = p.inbox0 ^ (step & 1);

        if (item == kItemFold) {
            // ---- tripole fold of the new velocities (fold_uv_kernel's work) ----
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
            if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
            if (tid < p.nranks) wait_counter<true>(p.topflag + tid, (unsigned long long)step + 1, p.error, p.timeout_ns);
            if (p.ll_on && tid >= 32 && tid < 34) {
                // the neighbours' row-ny values (before the fold) for my two ghost columns
                const int side = tid - 32;
                const unsigned long long *nbr = side == 0 ? p.wflag : p.eflag;
                double u, v;
                if (nbr && nbr[(a.ny - 1) / a.rows_per_chunk] != ~0ull &&
                    ll_take(p.ll_my, p.nyb, step, side, a.ny, u, v, p.error, p.timeout_ns)) {
                    const size_t at = (size_t)a.ny * ld + (side == 0 ? 0 : a.nxl + 1);
                    *(volatile double *)(u1 + at) = u;
                    *(volatile double *)(v1 + at) = v;
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(kTmaConsumers * 32) : "memory");
            const double *inbox = p.inbox[ibuf][p.rank];
            const int ild = a.inbox_ld;
            for (int c = tid; c <= a.nxl + 1; c += kTmaConsumers * 32) {
                int ig = a.lo + c;
                if (ig < 1) ig += a.nx_global;
                if (ig > a.nx_global) ig -= a.nx_global;
                const size_t top = (size_t)a.ny * ld + c;
                fold_necorner_column(u1, ld, a.ny, c, ig, a.nx_global, __ldcg(u1 + top), __ldcg(inbox + c),
                                     __ldcg(inbox + ild + c), -1.0, kFirstPair3D);
                fold_necorner_column(v1, ld, a.ny, c, ig, a.nx_global, __ldcg(v1 + top), __ldcg(inbox + 2 * ild + c),
                                     __ldcg(inbox + 3 * ild + c), -1.0, kFirstPair3D);
            }
            __threadfence();
            asm volatile("fence.proxy.async.global;" ::: "memory");
            asm volatile("bar.sync 1, %0;" ::"n"(kTmaConsumers * 32) : "memory");
            if (tid == 0) st_release_u64<false>(p.fold_flag, (unsigned long long)step + 1);
            continue;
        }

        const int strip = item % p.strips, chunk = item / p.strips;
        int r0, rend;
        chunk_rows(chunk, a.rows_per_chunk, a.ny, r0, rend);
        // (the last subcycle of evp, which also writes divu, shear, rdg_*, prs_sig, strint*, is a launch of
        // evp_subcycle_tma_kernel<true> after this kernel: a second instantiation of the sweep in here, inlined or
        // called, made the register allocator spill hundreds of bytes in the body that runs 119 times of 120)
        sweep_item<false, REVISED>(p, ring, full_bar, empty_bar, stage, phase, strip, r0, rend, par, ibuf, step, warp, lane);

        // ---- publish: this tile has finished subcycle `step` ----
        // Warps 1..3 only announce that their stores are issued (bar.arrive) and go on with the next item; warp 0
        // waits for the four warps (bar.sync), then ONE thread releases the counters.  The release is cumulative:
        // the stores of all 128 threads happen before the barrier completes, hence before the release fence of
        // lane 0, hence before the counter becomes visible (gpu scope; system scope where a peer GPU reads it or
        // the tile stored into peer memory).  A fence in every thread (first version) cost ~5 us of CTA time per item.
        if (warp != 0) {
            asm volatile("bar.arrive 2, %0;" ::"n"(kTmaConsumers * 32) : "memory");
        } else {
            // (barrier 2: barrier 1 belongs to the fold item, which warps 1..3 may reach while warp 0 is still here.
            // Warps 1..3 cannot arrive twice before warp 0 has been here once: every item spans more ring stages
            // than the ring holds, and a stage is recycled only when warp 0 has released it too.)
            asm volatile("bar.sync 2, %0;" ::"n"(kTmaConsumers * 32) : "memory");
            if (lane == 0) {
                const unsigned long long done = (unsigned long long)step + 1;
                const bool top = p.has_fold && rend >= a.ny;
                asm volatile("fence.proxy.async.global;" ::: "memory");
                if (!p.ll_on && p.nranks > 1) {
                    // (multi-rank without the mailbox: not a product configuration; kept for completeness)
                    st_release_u64<true>(p.tflag + item, done);
                    if (strip == 0 && p.pub_w) st_release_u64<true>(p.pub_w + chunk, done);
                    if (strip == p.strips - 1 && p.pub_e) st_release_u64<true>(p.pub_e + chunk, done);
                } else {
                    st_release_u64<false>(p.tflag + item, done);
                }
                if (top) {
                    // top tiles stored into the tripole inboxes of other ranks (peer memory).  Each one releases at
                    // gpu scope (above) before it counts itself; the one that completes the count takes a SYSTEM fence
                    // and tells every rank -- release is cumulative, so all top tiles' inbox stores are covered: one
                    // system-scope fence per rank and subcycle.
                    const unsigned long long ntop = (unsigned long long)p.item_count[2];
                    const unsigned long long n = atomicAdd(p.top_count, 1ull) + 1ull;
                    if (n == ntop * done) {
                        __threadfence_system();
                        for (int r = 0; r < p.nranks; ++r) st_release_u64<true>(p.pub_top[r] + p.rank, done);
                    }
                }
            }
            __syncwarp();
        }
    }
}

}  // namespace evp
