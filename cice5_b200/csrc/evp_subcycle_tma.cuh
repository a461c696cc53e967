// evp_subcycle_tma.cuh -- fused subcycle kernel v2: persistent, warp-specialised, TMA-fed.
//
// Same arithmetic and the same warp-sweep as evp_subcycle_kernel (evp_kernels.cuh; reference
// stress source/ice_dyn_evp.F90:618-847 and stepu source/ice_dyn_shared.F90:700-746), but the 36
// input planes of a row are no longer loaded into registers by the computing warps.  A producer
// warp streams them into a shared-memory ring with TMA (cp.async.bulk.tensor over ONE 3-D tensor
// map of the whole plane arena: box = 128 columns x 1 row x k consecutive planes), completion is
// signalled through mbarriers, and the consumer warps only see ~30-cycle shared-memory reads.
// Bytes in flight per SM are set by the ring depth, not by how many warps fit at 128 registers.
//
// CTA = 4 consumer warps (31 U columns each -> 124 U columns, box column 0 = the west ghost of
// the strip) + 1 producer warp.  Persistent: CTA b takes work items b, b + grid, ... ; an item is
// (strip, row chunk).  Stage j of an item holds: state planes (u, v, 12 stresses) of row j,
// T coefficients + activity mask of row j, U coefficients of row j-1.
#pragma once
#include <cuda.h>
#include <cstdio>

#include "evp_kernels.cuh"

namespace evp {

#ifndef EVP_TMA_WARPS
#define EVP_TMA_WARPS 4
#endif
constexpr int kTmaConsumers = EVP_TMA_WARPS;                 // 4 (128-column boxes) or 8 (256-column boxes)
constexpr int kTmaBoxW = 32 * kTmaConsumers;                 // doubles per box row
static_assert(kTmaBoxW <= 256, "TMA box dimensions are limited to 256 elements");
constexpr int kTmaUCols = kTmaConsumers * kUColsPerWarp;     // 124
#ifndef EVP_TMA_STAGES
#define EVP_TMA_STAGES 2
#endif
#ifndef EVP_TMA_BLOCKS
#define EVP_TMA_BLOCKS 3
#endif
#ifndef EVP_TMA_LATE
#define EVP_TMA_LATE 1
#endif
// EVP_TMA_LATE = 1: the U coefficients are read from the stage just before stepu and the stage is
// released after the row is finished (fewer live registers, stage held longer).
constexpr bool kLateRelease = EVP_TMA_LATE != 0;
constexpr int kTmaStages = EVP_TMA_STAGES;
constexpr int kTmaBlocksPerSM = EVP_TMA_BLOCKS;
constexpr int kStateDepth = 14, kTDepth = 12, kUDepth = 10;
// T slots of a stage (each kTmaBoxW doubles).  Compact row: 0 HTE, 1 HTN, 2 strength, 3 tinyarea;
// full row: 0 strength, 1 tinyarea, 2..9 dxt dyt dxhy dyhx cxp cyp cxm cym; both: 10 tarear (last
// subcycle), 11 HTN of this row when only the next row needs it (its southern HTN).
constexpr int kSlotTarear = 10, kSlotHtnExtra = 11;
constexpr int kSortThreads = 64;
constexpr int kItemBoundary = 0x40000000;    // tag of a work item that stores into peer memory
// stage word bits written by the producer
constexpr unsigned kMetaState = 1u, kMetaCompact = 2u, kMetaTinyFromMetrics = 4u, kMetaHtnExtra = 8u, kMetaT = 16u;
constexpr int kMaskBoxW = 128 + 16;          // per mask box; kTmaBoxW / 128 boxes per row
constexpr int kMaskBoxes = kTmaBoxW / 128;
constexpr int kMaskFlagOff = 256 * (kMaskBoxes - 1) + kMaskBoxW + 48;     // stage word after the mask bytes (16-byte aligned)     // the byte mask is fetched from a 16-byte aligned column (TMA
                                             // needs 16-byte aligned box starts), up to 12 bytes early
constexpr int kStageBytes = (kStateDepth + kTDepth + kUDepth) * kTmaBoxW * 8 + 256 * kMaskBoxes;   // + mask row
static_assert(kStageBytes % 128 == 0, "TMA destinations must stay 128-byte aligned");

struct TmaArgs {
    LoopArgs a;                  // geometry, constants, output / push pointers (inputs unused)
    int plane_state;             // first plane of the OLD state copy
    int plane_t, plane_u;        // first plane of the T / U coefficient groups
    int strips, chunks, items;
    // dynamic work queue: CTAs draw positions with atomicAdd(queue, 1); position q maps to the item
    // item_list[q] (the items of this call that hold ice, built by item_list_kernel) for
    // q < *item_count.  One counter per subcycle launch, zeroed once per evp call.
    int *queue;
    const int *item_list, *item_count;
    const unsigned char *tile_flags;   // [strip][ny + 4]: bit0 any active T cell, bit1 any active U cell
    int flags_ld;
    int water_is_ocn;            // cosw == 1 && sinw == 0: waterx/watery are not fetched (== uocn/vocn)
    double puny;                 // tinyarea = puny*tarea (ice_grid.F90:354)
    // fused cross-rank signal: every CTA counts itself in `done` when its stores are out; the last
    // one bumps the epoch and releases it to the peers (what halo_signal_kernel does otherwise)
    int fused_signal;
    int *done, *done_bnd;        // finished CTAs / finished boundary items of this launch
    SyncArgs sync;
    // timing trace of this launch (EVP_B200_TRACE=1), globaltimer ns: [0] first CTA start, [1] last
    // CTA done, [2] peers signalled; the fold kernel adds [3] wait begin, [4] wait end
    unsigned long long *trace;
};

// Rows of a chunk: T rows r0 .. rend, U rows r0 .. rend-1
__host__ __device__ __forceinline__ void chunk_rows(int chunk, int rpc, int ny, int &r0, int &rend)
{
    r0 = 1 + chunk * rpc;
    rend = r0 + rpc < ny + 1 ? r0 + rpc : ny + 1;
}

__device__ __forceinline__ unsigned smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *b, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *b, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *b)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
// try_wait with a suspend-time hint: the waiting thread is parked by the hardware until the phase completes (or
// the hint, in ns, runs out) instead of spinning -- in the first loop-kernel profile the producers' spin on
// `empty` (SYNCS.TRYWAIT + YIELD + BRA, 2.8e9 rounds per call) was a quarter of all executed warp instructions,
// issue slots taken from the FP64 consumers on the same schedulers
#ifndef EVP_TMA_WAIT_HINT_NS
#define EVP_TMA_WAIT_HINT_NS 20000
#endif
__device__ __forceinline__ void mbar_wait(unsigned long long *b, unsigned parity)
{
    unsigned done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(b)), "r"(parity), "r"((unsigned)EVP_TMA_WAIT_HINT_NS)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, unsigned long long *bar, int x,
                                            int y, int z)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, unsigned long long *bar, int x, int y)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y)
        : "memory");
}

// dxt dyt dxhy dyhx cxp cyp cxm cym from HTE(i,j), HTE(i-1,j), HTN(i,j), HTN(i,j-1) exactly as the
// reference builds them (ice_grid.F90:356-357,361-366,1456,1536)
__device__ __forceinline__ void metric_terms(double hte, double hte_w, double htn, double htn_s, double *m)
{
    m[0] = __dmul_rn(0.5, __dadd_rn(htn, htn_s));                               // dxt
    m[1] = __dmul_rn(0.5, __dadd_rn(hte, hte_w));                               // dyt
    m[2] = __dmul_rn(0.5, __dsub_rn(hte, hte_w));                               // dxhy
    m[3] = __dmul_rn(0.5, __dsub_rn(htn, htn_s));                               // dyhx
    m[4] = __dsub_rn(__dmul_rn(1.5, htn), __dmul_rn(0.5, htn_s));               // cxp
    m[5] = __dsub_rn(__dmul_rn(1.5, hte), __dmul_rn(0.5, hte_w));               // cyp
    m[6] = -__dsub_rn(__dmul_rn(1.5, htn_s), __dmul_rn(0.5, htn));              // cxm
    m[7] = -__dsub_rn(__dmul_rn(1.5, hte_w), __dmul_rn(0.5, hte));              // cym
}

// Consumer warps of one CTA: `n` more boundary items are complete (all their stores into peer
// memory issued).  Fence once, count; the CTA that completes the count releases the epoch.
__device__ __noinline__ void boundary_done(const TmaArgs &p, int *bnd_count, int warp, int lane)
{
    __threadfence_system();                          // my stores into peer memory are out
    asm volatile("bar.sync 1, %0;" ::"n"(kTmaConsumers * 32) : "memory");
    if (warp == 0) {
        int fire = 0;
        if (lane == 0) {
            const int n = *bnd_count;                // written by this thread only
            *bnd_count = 0;
            __threadfence_system();
            fire = atomicAdd(p.done_bnd, n) + n == p.item_count[1];
            if (fire) __threadfence_system();
        }
        fire = __shfl_sync(0xffffffffu, fire, 0);
        if (fire) {
            peer_signal_warp(p.sync, lane);
            if (p.trace && lane == 0) p.trace[2] = global_ns();
        }
    }
}

template <bool LAST, bool REVISED = false>
__global__ void __launch_bounds__((kTmaConsumers + 1) * 32, kTmaBlocksPerSM)
evp_subcycle_tma_kernel(const __grid_constant__ CUtensorMap mapS, const __grid_constant__ CUtensorMap mapS2,
                        const __grid_constant__ CUtensorMap mapT1,
                        const __grid_constant__ CUtensorMap mapT3, const __grid_constant__ CUtensorMap mapT4,
                        const __grid_constant__ CUtensorMap mapT10, const __grid_constant__ CUtensorMap mapU,
                        const __grid_constant__ CUtensorMap mapM, const __grid_constant__ TmaArgs p)
{
    extern __shared__ __align__(1024) unsigned char ring[];
    __shared__ unsigned long long full_bar[kTmaStages], empty_bar[kTmaStages];
    __shared__ int bnd_count;                        // boundary items taken by this CTA (thread 0 only)
    const LoopArgs &a = p.a;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        bnd_count = 0;
#pragma unroll
        for (int s = 0; s < kTmaStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], kTmaConsumers); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    int stage = 0;
    unsigned phase = 0;

    if (warp == kTmaConsumers) {
        // ------------------------------ producer ------------------------------
        if (lane == 0) {
            const int nitems = p.item_count[0], nbnd = p.item_count[1];
            for (;;) {
                const int q = atomicAdd(p.queue, 1);
                if (q >= nitems) {                                      // queue drained: tell the consumers
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    unsigned char *mk = ring + (size_t)stage * kStageBytes + (kStateDepth + kTDepth + kUDepth) * kTmaBoxW * 8;
                    *reinterpret_cast<volatile int *>(mk + kMaskFlagOff + 4) = -1;
                    mbar_arrive(&full_bar[stage]);
                    break;
                }
                const int item = p.item_list[q];
                const int item_tag = item | (q < nbnd ? kItemBoundary : 0);
                const int strip = item % p.strips, chunk = item / p.strips;
                const int x0 = strip * kTmaUCols;                       // box column 0 = U column c0 - 1
                int r0, rend;
                chunk_rows(chunk, a.rows_per_chunk, a.ny, r0, rend);
                const unsigned char *fl = p.tile_flags + (size_t)strip * p.flags_ld;
                for (int j = r0 - 1; j <= rend; ++j) {
                    // what this row segment needs: T data if it has an active T cell; U data of the
                    // row below if that row has an active U cell; the state (u, v, stresses) if this
                    // row or the next one has an active T cell or this row an active U cell.  The top
                    // rows are never skipped: their velocities feed the tripole inboxes every subcycle.
                    const bool top = j >= a.ny - 2;
                    const unsigned f0 = fl[j], f1 = fl[j + 1], fm1 = j > 0 ? fl[j - 1] : 0u;
                    const bool needT = j >= r0 && ((f0 & 1u) || top);
                    const bool needU = j > r0 && ((fm1 & 2u) || top);
                    const bool needS = ((f0 | f1) & 1u) || (f0 & 2u) || top;
                    const bool cmp = (f0 & 4u) != 0u, tinym = (f0 & 8u) != 0u;
                    // the next row, if compact, recomputes its metrics from HTN of this row too
                    const bool nextT = j + 1 <= rend && ((f1 & 1u) || j + 1 >= a.ny - 2) && (f1 & 4u);
                    const bool htn_extra = nextT && !(needT && cmp);
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    unsigned char *sb = ring + (size_t)stage * kStageBytes;
                    unsigned char *tb = sb + kStateDepth * kTmaBoxW * 8;
                    unsigned char *mk = sb + (kStateDepth + kTDepth + kUDepth) * kTmaBoxW * 8;
                    *reinterpret_cast<volatile unsigned *>(mk + kMaskFlagOff) =
                        (needS ? kMetaState : 0u) | (cmp ? kMetaCompact : 0u) | (tinym ? kMetaTinyFromMetrics : 0u) |
                        (htn_extra ? kMetaHtnExtra : 0u) | (needT ? kMetaT : 0u);
                    *reinterpret_cast<volatile int *>(mk + kMaskFlagOff + 4) = item_tag;
                    const int tdepth = cmp ? (tinym ? 3 : 4) : 10;
                    unsigned bytes = kMaskBoxW * kMaskBoxes;
                    const bool uv_only = j == r0 - 1;                    // first row of an item: velocities only
                    if (needS) bytes += (uv_only ? 2 : kStateDepth) * kTmaBoxW * 8;
                    if (needT) bytes += (tdepth + (LAST ? 1 : 0)) * kTmaBoxW * 8;
                    if (htn_extra) bytes += kTmaBoxW * 8;
                    if (needU) bytes += (p.water_is_ocn ? kUDepth - 2 : kUDepth) * kTmaBoxW * 8;
                    mbar_expect_tx(&full_bar[stage], bytes);
#pragma unroll
                    for (int mb = 0; mb < kMaskBoxes; ++mb)
                        tma_load_2d(mk + mb * 256, &mapM, &full_bar[stage], (x0 & ~15) + mb * 128, j);
                    if (needS) tma_load_3d(sb, uv_only ? &mapS2 : &mapS, &full_bar[stage], x0, j, p.plane_state);
                    if (needT) {
                        if (!cmp) tma_load_3d(tb, &mapT10, &full_bar[stage], x0, j, p.plane_t + 2);
                        else if (tinym) tma_load_3d(tb, &mapT3, &full_bar[stage], x0, j, p.plane_t);
                        else tma_load_3d(tb, &mapT4, &full_bar[stage], x0, j, p.plane_t);
                        if (LAST)
                            tma_load_3d(tb + kSlotTarear * kTmaBoxW * 8, &mapT1, &full_bar[stage], x0, j, p.plane_t + 12);
                    }
                    if (htn_extra)
                        tma_load_3d(tb + kSlotHtnExtra * kTmaBoxW * 8, &mapT1, &full_bar[stage], x0, j, p.plane_t + 1);
                    if (needU)
                        tma_load_3d(sb + (kStateDepth + kTDepth) * kTmaBoxW * 8, &mapU, &full_bar[stage], x0, j - 1,
                                    p.plane_u);
                    if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
                }
            }
        }
        return;
    }

    // ------------------------------ consumers ------------------------------
    const Consts &k = a.k;
    const int ld = a.ld;
    const int bi = 1 + warp * kUColsPerWarp + lane;                     // my column inside the box
    bool in_bnd = false;                                                // boundary items taken, not yet counted
    if (p.trace && blockIdx.x == 0 && threadIdx.x == 0) p.trace[0] = global_ns();
    for (;;) {
        mbar_wait(&full_bar[stage], phase);                             // first stage of the next item
        const int item_tag = *reinterpret_cast<const int *>(
            ring + (size_t)stage * kStageBytes + (kStateDepth + kTDepth + kUDepth) * kTmaBoxW * 8 + kMaskFlagOff + 4);
        if (in_bnd && (item_tag < 0 || !(item_tag & kItemBoundary))) {
            // this CTA has left the boundary items (they come first in the queue): count them;
            // whoever completes the count signals the peers, and the interior items of every rank
            // are computed while the signals travel
            if (p.fused_signal) boundary_done(p, &bnd_count, warp, lane);
            in_bnd = false;
        }
        if (item_tag < 0) break;
        const int item = item_tag & ~kItemBoundary;
        if (item_tag & kItemBoundary) {
            in_bnd = true;
            if (threadIdx.x == 0) ++bnd_count;
        }
        const int strip = item % p.strips, chunk = item / p.strips;
        const int c = 1 + strip * kTmaUCols + warp * kUColsPerWarp + lane;   // T column (U column if lane < 31)
        const bool warp_on = (c - lane) <= a.nxl;                       // warp-uniform
        const int mshift = (strip * kTmaUCols) & 15;                    // mask box starts this many bytes early
        int r0, rend;
        chunk_rows(chunk, a.rows_per_chunk, a.ny, r0, rend);
        const bool tcol = c <= a.nxl + 1;
        const bool ucol = lane < kUColsPerWarp && c <= a.nxl;
        const bool tstore_col = tcol && (lane < kUColsPerWarp || c == a.nxl + 1);

        // row r0-1: velocities only
        double u_s, v_s, uw_s, vw_s;
        double htn_s = 0.0;                                             // HTN of the row below (compact rows)
        bool loaded_s;                                                  // state of the row below was fetched
        {
            const double *st = reinterpret_cast<const double *>(ring + (size_t)stage * kStageBytes);
            const unsigned meta = *reinterpret_cast<const unsigned *>(
                ring + (size_t)stage * kStageBytes + (kStateDepth + kTDepth + kUDepth) * kTmaBoxW * 8 + kMaskFlagOff);
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
            const bool any_u = __ballot_sync(0xffffffffu, uact) != 0u;
            double s[12], tco[11], uco[kUDepth];
            // tco: dxt dyt dxhy dyhx cxp cyp cxm cym tinyarea strength tarear
            if (any_t) {
#pragma unroll
                for (int q = 0; q < 12; ++q) s[q] = st[(2 + q) * kTmaBoxW + bi];
                if (meta & kMetaCompact) {
                    // metric terms from HTE (this cell and its western neighbour) and HTN (this
                    // row and the row below), the reference's own formulas (ice_grid.F90:354-367,
                    // 1456,1536); set_grid checked that they reproduce the host arrays bit for bit
                    // on this row segment.  Explicit _rn intrinsics: never contracted.
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
            if (!kLateRelease) {
                if (any_u) {
#pragma unroll
                    for (int q = 0; q < kUDepth - 2; ++q) uco[q] = uc[q * kTmaBoxW + bi];
                    uco[8] = p.water_is_ocn ? uco[1] : uc[8 * kTmaBoxW + bi];
                    uco[9] = p.water_is_ocn ? uco[2] : uc[9 * kTmaBoxW + bi];
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty_bar[stage]);          // slot free: everything is in registers
                if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
            }

            const size_t idx = (size_t)r * ld + c;
            double str[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            if (tact) {
                double divu, shear, rdg_conv, rdg_shear, prs_sig;
#ifdef EVP_TMA_NOCOMPUTE   /* timing experiment only: data path without the arithmetic */
                for (int q = 0; q < 8; ++q) str[q] = s[q] + tco[q] + u_c + uw_c + v_c + vw_c; divu = shear = rdg_conv = rdg_shear = prs_sig = 0.0;
#else
                stress_cell<LAST>(k, u_c, uw_c, u_s, uw_s, v_c, vw_c, v_s, vw_s, tco[0], tco[1], tco[2], tco[3],
                                  tco[4], tco[5], tco[6], tco[7], tco[8], tco[9], LAST ? tco[10] : 0.0, s,
                                  str, divu, shear, rdg_conv, rdg_shear, prs_sig);
#endif
                if (tstore_col && (r < rend || r == a.ny + 1)) {
#pragma unroll
                    for (int q = 0; q < 12; ++q) a.s1[q][idx] = s[q];
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
                    if (kLateRelease) {
#pragma unroll
                        for (int q = 0; q < kUDepth - 2; ++q) uco[q] = uc[q * kTmaBoxW + bi];
                        uco[8] = p.water_is_ocn ? uco[1] : uc[8 * kTmaBoxW + bi];
                        uco[9] = p.water_is_ocn ? uco[2] : uc[9 * kTmaBoxW + bi];
                    }
                    // uco: aiu*rhow*Cw uocn vocn forcex forcey umassdti fm uarear waterx watery
                    const double uold = u_s, vold = v_s;
                    const double du = uco[1] - uold, dv = uco[2] - vold;
#ifdef EVP_TMA_NOCOMPUTE
                    const double vrel = uco[0] + du + dv;
#else
                    const double vrel = uco[0] * sqrt(du * du + dv * dv);
#endif
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
#ifdef EVP_TMA_NOCOMPUTE
                    unew = cca + cc1 + ab2; vnew = ccb + cc2;
#else
                    unew = (cca * cc1 + ccb * cc2) / ab2;
                    vnew = (cca * cc2 - ccb * cc1) / ab2;
#endif
                    if (LAST) { a.strintx[iu] = strintx; a.strinty[iu] = strinty; }
                }
                if (ucol) {
                    a.u1[iu] = unew; a.v1[iu] = vnew;
                    if (c == 1 && a.uW) {
                        const size_t ig = (size_t)(r - 1) * ld + a.nxl_west + 1;
                        a.uW[ig] = unew; a.vW[ig] = vnew;
                    }
                    if (c == a.nxl && a.uE) {
                        const size_t ig = (size_t)(r - 1) * ld;
                        a.uE[ig] = unew; a.vE[ig] = vnew;
                    }
                    if (a.tripole && r - 1 >= a.ny - 1) {
                        tripole_push(a, c, (r - 1) - (a.ny - 1), unew, vnew);
                       
                    }
                }
            }
            if (kLateRelease) {
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty_bar[stage]);          // row done: slot free
                if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
            }
            s1p = str[0]; s5p = str[4]; n2p = e2; n7p = e7;
            u_s = u_c; v_s = v_c; uw_s = uw_c; vw_s = vw_c; mprev = m; loaded_s = loaded_c;
        }
    }
    if (p.fused_signal || p.trace) {
        __threadfence_system();
        asm volatile("bar.sync 1, %0;" ::"n"(kTmaConsumers * 32) : "memory");
        if (warp == 0) {
            int last = 0;
            if (lane == 0) {
                __threadfence_system();
                last = atomicAdd(p.done, 1) == (int)gridDim.x - 1;
                if (last) {
                    __threadfence_system();
                    if (p.trace) p.trace[1] = global_ns();
                }
            }
            last = __shfl_sync(0xffffffffu, last, 0);
            if (last && p.fused_signal && p.item_count[1] == 0) {       // no boundary item held ice
                peer_signal_warp(p.sync, lane);
                if (p.trace && lane == 0) p.trace[2] = global_ns();
            }
        }
    }
}

// Row activity of every strip, once per evp call (the masks are fixed for the call): lets the
// producer skip the TMA boxes of row segments without ice (land, open ocean).
__global__ void tile_flags_kernel(const unsigned char *mask, int ldm, int nxl, int ny, int strips,
                                  unsigned char *flags, int flags_ld, const unsigned char *gflags)
{
    const int strip = blockIdx.x, r = blockIdx.y * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r > ny + 3) return;
    unsigned f = 0;
    if (r >= 1 && r <= ny + 1) {
        const int c0 = 1 + strip * kTmaUCols;
        for (int q = lane; q <= kTmaUCols; q += 32) {          // T columns c0 .. c0+124
            const int c = c0 + q;
            if (c <= nxl + 1) {
                const unsigned m = mask[(size_t)r * ldm + c];
                f |= (m & 1u);
                if (q < kTmaUCols && c <= nxl) f |= (m & 2u);
            }
        }
    }
    f = __reduce_or_sync(0xffffffffu, f);
    if (lane == 0)
        flags[(size_t)strip * flags_ld + r] = (unsigned char)(f | (gflags[(size_t)strip * flags_ld + r] & 12u));
}

// Once per set_grid: which row segments of which strips can use compact metric rows.  bit2: the 8
// metric terms of every T cell of the segment equal metric_terms(HTE, HTN) bit for bit; bit3: so
// does tinyarea = puny*(dxt*dyt).  All cells are checked, ocean or not.
struct MetricCheckArgs {
    const double *hte, *htn, *met[8], *tinyarea;
    int ld, nxl, ny;
    unsigned char *flags;
    int flags_ld;
};
__global__ void metric_flags_kernel(const __grid_constant__ MetricCheckArgs a)
{
    const int strip = blockIdx.x, r = blockIdx.y * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r > a.ny + 3) return;
    unsigned bad = 0;                                          // bit0: a metric differs, bit1: tinyarea differs
    if (r >= 1 && r <= a.ny + 1) {
        const int c0 = 1 + strip * kTmaUCols;
        for (int q = lane; q <= kTmaUCols; q += 32) {
            const int c = c0 + q;
            if (c > a.nxl + 1) continue;
            const size_t i = (size_t)r * a.ld + c;
            double m[8];
            metric_terms(a.hte[i], a.hte[i - 1], a.htn[i], a.htn[i - a.ld], m);
#pragma unroll
            for (int t = 0; t < 8; ++t)
                if (__double_as_longlong(m[t]) != __double_as_longlong(a.met[t][i])) bad |= 1u;
            const double tiny = __dmul_rn(1.0e-11, __dmul_rn(m[0], m[1]));
            if (__double_as_longlong(tiny) != __double_as_longlong(a.tinyarea[i])) bad |= 2u;
        }
    } else {
        bad = 3u;
    }
    bad = __reduce_or_sync(0xffffffffu, bad);
    if (lane == 0)
        a.flags[(size_t)strip * a.flags_ld + r] = (unsigned char)((bad & 1u) ? 0u : ((bad & 2u) ? 4u : 12u));
}

// The work list of the call: the (strip, row chunk) items that hold ice, heaviest first (most rows
// with ice: longest-processing-time order, so the dynamic queue ends evenly even with 2-3 items
// per CTA).  An item without any active row segment writes nothing (the ping-pong copies agree
// where the loop never writes), so it is dropped -- except the chunks that reach the top rows of a
// tripole grid, whose velocities feed the fold every subcycle.  The order inside one weight class
// is not deterministic; results do not depend on the order of the items.
// Per call, for the persistent loop kernel (evp_loop_tma.cuh): `natural` keeps the items in item order (south to
// north, the order the dependency wavefront wants) and the dependency counters are reset -- 0 for an item
// that will run, all ones for an item without ice (it never runs, and never has to be waited for); the edge
// strips' counters also into the E-W neighbours' ghost arrays.  count[2] = items that reach the top rows.
struct ItemListLoop {
    int natural;
    unsigned long long *misc;              // topflag[8], fold_flag, top_count: zeroed
    unsigned long long *tflag, *pub_w, *pub_e;
};
__global__ void item_list_kernel(const unsigned char *flags, int flags_ld, int strips, int chunks, int rpc,
                                 int ny, int tripole, int bnd_west, int bnd_east, int *list, int *count, int *queues,
                                 int nqueues, ItemListLoop lp)
{
    // Stable counting sort, one block of kSortThreads threads, thread t owns a contiguous range of
    // items.  key = weight (rows of the item with ice, 0 .. rpc + 3 <= 67) + kW for boundary items
    // (they store into peer memory: E-W ghost columns, tripole inboxes): boundary items first, so
    // that the peers can be signalled while the interior items are still being computed, heaviest
    // first inside each class; equal keys keep the item order (neighbouring items run at the same
    // time and the list is the same on every call).  count[0] = items kept, count[1] = boundary.
    constexpr int kW = 80, kBins = 2 * kW;
    __shared__ int hist[kSortThreads][kBins];
    __shared__ int binbase[kBins];
    __shared__ int ntop;
    const int t = threadIdx.x;
    if (t == 0) ntop = 0;
    if (lp.misc && t < 10) lp.misc[t] = 0ull;
    for (int q = t; q < nqueues; q += blockDim.x) queues[q] = 0;
    for (int q = 0; q < kBins; ++q) hist[t][q] = 0;
    __syncthreads();
    const int items = strips * chunks;
    const int per = (items + kSortThreads - 1) / kSortThreads;
    const int i0 = min(t * per, items), i1 = min(i0 + per, items);
    for (int pass = 0; pass < 2; ++pass) {
        for (int item = i0; item < i1; ++item) {
            const int strip = item % strips, chunk = item / strips;
            int r0, rend;
            chunk_rows(chunk, rpc, ny, r0, rend);
            const unsigned char *fl = flags + (size_t)strip * flags_ld;
            int w = 0;
            for (int j = r0 - 1; j <= rend + 1; ++j) w += (fl[j] & 3u) != 0u;    // activity bits only
            if (w == 0 && tripole && rend >= ny - 2) w = 1;
            if (pass == 0 && lp.tflag) {
                const unsigned long long v = w == 0 ? ~0ull : 0ull;
                lp.tflag[item] = v;
                if (strip == 0 && lp.pub_w) lp.pub_w[chunk] = v;
                if (strip == strips - 1 && lp.pub_e) lp.pub_e[chunk] = v;
                if (w != 0 && tripole && rend >= ny) atomicAdd(&ntop, 1);
            }
            if (w == 0) continue;
            const bool bnd = (strip == 0 && bnd_west) || (strip == strips - 1 && bnd_east) || (tripole && rend >= ny - 2);
            const int key = lp.natural ? 1 : min(w, kW - 1) + (bnd ? kW : 0);
            if (pass == 0) ++hist[t][key];
            else list[binbase[key] + hist[t][key]++] = item;
        }
        __syncthreads();
        if (pass == 0) {
            if (t == 0) {
                int off = 0;
                count[1] = 0;
                for (int key = kBins - 1; key >= 1; --key) {
                    if (key == kW - 1) count[1] = off;          // everything before: boundary items
                    binbase[key] = off;
                    for (int q = 0; q < kSortThreads; ++q) {    // exclusive scan over the threads
                        const int n = hist[q][key];
                        hist[q][key] = off - binbase[key];
                        off += n;
                    }
                }
                count[0] = off;
                count[2] = ntop;
            }
            __syncthreads();
        }
    }
}

}  // namespace evp
