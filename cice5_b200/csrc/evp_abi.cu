// evp_abi.cu -- host side of the C ABI in include/evp_b200.h: the B200-native replacement of
// CICE5's `evp(dt)` driver (source/ice_dyn_evp.F90:68-510).  All arithmetic runs in the CUDA
// kernels of evp_kernels.cuh / evp_prep.cuh; there is no CPU fallback.
#include "../../include/evp_b200.h"

#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <new>
#include <vector>

#include "evp_kernels.cuh"
#include "evp_prep.cuh"
#include "evp_subcycle_tma.cuh"
#include "evp_loop_tma.cuh"
#include "evp_strength.cuh"
#include "evp_remap.cuh"
#include "evp_eap.cuh"

using namespace evp;

namespace {

thread_local char g_cuda_err[512] = "";

#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            snprintf(g_cuda_err, sizeof(g_cuda_err), "%s:%d %s: %s", __FILE__, __LINE__, #call, \
                     cudaGetErrorString(e_));                                                 \
            return EVP_ERR_CUDA;                                                              \
        }                                                                                     \
    } while (0)
#define TRY(x) do { int rc_ = (x); if (rc_ != EVP_OK) return rc_; } while (0)
// inside evp_gpu_init, once the handle exists: a failing call releases everything the handle holds so far
#define CKH(call)                                                                             \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            snprintf(g_cuda_err, sizeof(g_cuda_err), "%s:%d %s: %s", __FILE__, __LINE__, #call, \
                     cudaGetErrorString(e_));                                                 \
            evp_gpu_finalize(h);                                                              \
            return EVP_ERR_CUDA;                                                              \
        }                                                                                     \
    } while (0)

// ---- plane directory ----------------------------------------------------------------------
enum Plane {
    // T-cell coefficients of the subcycle kernel (TMA boxes over consecutive planes):
    //   compact rows: HTE, HTN, strength [, tinyarea]  (depth 3 / 4 from G_HTE) -- the 8 metric
    //                 terms (and tinyarea) are recomputed from HTE / HTN where set_grid verified
    //                 that this reproduces the host's arrays bit for bit
    //   full rows   : strength, tinyarea, dxt .. cym   (depth 10 from W_STRENGTH)
    //   tarear alone on the last subcycle, HTN alone where the next row needs its southern HTN
    G_HTE, G_HTN, W_STRENGTH, G_TINYAREA, G_DXT, G_DYT, G_DXHY, G_DYHX, G_CXP, G_CYP, G_CXM, G_CYM, G_TAREAR,
    // U-cell coefficients of the subcycle kernel: contiguous, one TMA box of depth 10
    // (waterx, watery last: when cosw = 1, sinw = 0 they equal uocn, vocn and the box stops before them)
    W_VRELC, I_UOCN, I_VOCN, W_FORCEX, W_FORCEY, W_UMASSDTI, S_FM, G_UAREAR, W_WATERX, W_WATERY,
    // dynamic state, ping-pong copies A / B: u, v, stressp_1..4, stressm_1..4, stress12_1..4
    S_UA, S_VA, S_STRA, S_STRA_END = S_STRA + 11,
    S_UB, S_VB, S_STRB, S_STRB_END = S_STRB + 11,
    // remaining grid planes and per-call inputs
    G_TAREA, G_UAREA, G_FCOR,
    I_AICE, I_VICE, I_VSNO, I_AICE_INIT, I_STRAIRXT, I_STRAIRYT, I_SSTLTX, I_SSTLTY, I_STRENGTH,
    // evp locals (ice_dyn_evp.F90:120-137)
    W_TMASS, W_UMASS, W_ICETMASK, W_1, W_2, W_AIU, I_CW,
    // outputs (evp_out_t order, first 17; contiguous: zeroed with one memset)
    O_STRAIRX, O_STRAIRY, O_STRTLTX, O_STRTLTY, O_STROCNX, O_STROCNY, O_STRINTX, O_STRINTY,
    O_STROCNXT, O_STROCNYT, O_DIVU, O_SHEAR, O_RDGCONV, O_RDGSHEAR, O_PRSSIG, O_UVELINIT, O_VVELINIT,
    NPLANES
};
constexpr int kTGroup = G_HTE, kUGroup = W_VRELC;

constexpr int kEapTablePitch = (kEapTableElems + 31) / 32 * 32;
constexpr int kEapPlanes = 51;   // 8 tensor + 2 averages + 9 diagnostics + 16 rotation + 12 yield stresses + 4 alphar

struct PeerBlob {                 // exchanged between the tasks of one node
    cudaIpcMemHandle_t arena;
    int rank, device;
    long long pid;                // same pid => same process: use the raw pointer, no IPC
    unsigned long long arena_ptr;
};

struct Launcher;
struct Op {                       // one step of the evp(dt) program of a handle
    bool sync;                    // true: cross-rank barrier (signal + wait), no fn
    std::function<int(Launcher &)> fn;
};

}  // namespace

struct evp_handle {
    evp_config_t cfg;
    Geom g;
    InboxGeom ib;
    Consts k;
    int nxb, nyb, ld, ldm;
    size_t plane_elems;           // doubles per plane (ld * nyb)
    // one arena, identical layout on every rank, so a peer's plane is peer_base + offset
    char *arena = nullptr;
    size_t arena_bytes = 0;
    size_t off_planes = 0, off_inbox = 0, off_flags = 0, off_int = 0, off_mask = 0, off_counts = 0, off_tflags = 0;
    int tflags_ld = 0, tstrips = 0;
    std::vector<void *> pinned;                  // host arrays page-locked by evp_gpu_pin_host_arrays
    // dense staging area for host <-> device copies (kRepackMax arrays of nxb * nyb doubles), allocated on first use
    char *stage = nullptr;
    std::vector<std::pair<void *, int>> batch_dst;   // pending batch: (plane pointer, is_int) in staging order
    // resident-state protocol of evp_gpu_run (evp_gpu_set_resident)
    bool resident = false, state_valid = false;
    unsigned dl_mask = EVP_DL_ALL;
    bool strair_given = false;                   // evp_in_t.strax/stray given: strairx/y are copies of them
    int t_rpc = 0, t_chunks = 0, t_items = 0;    // TMA kernel work items: (strip, chunk of t_rpc rows)
    size_t off_trace = 0;                        // u64 trace[ndte][8] (EVP_B200_TRACE=1)
    bool trace_on = false;
    size_t off_gflags = 0;                       // per strip row: bit2 metrics, bit3 tinyarea reproducible from HTE/HTN
    size_t off_queue = 0;                        // int queue[ndte], done[ndte], done_bnd[ndte], item_count[4], item_list[t_items]
    // dependency counters of the persistent loop kernel (evp_loop_tma.cuh), u64: topflag[8], fold_flag, top_count,
    // pad to 16, wflag[t_chunks], eflag[t_chunks], tflag[t_items].  Peers write topflag / wflag / eflag, so the
    // region sits before every per-rank-sized one and t_chunks is the same on all ranks.
    size_t off_lflags = 0;
    size_t off_inbox_wide = 0;    // 2 buffers x kWideSlots slots of the inbox layout: halos of plane stacks (transport)
    size_t off_ll = 0;            // ghost mailbox of the loop kernel: u64[kLLBufs][2][nyb][4], written by the E-W neighbours
    bool use_loop = false;        // the ndte loop as ONE persistent kernel (default with the TMA kernel; EVP_B200_LOOP=0: off)
    char *peer_arena[kMaxRanks] = {nullptr};
    bool peer_opened[kMaxRanks] = {false};
    bool grid_set = false, uploaded = false, peers_set = false;
    int parity = 0;               // which ping-pong copy holds the current state (0 = A)
    int halo_seq = 0;             // halo operations issued in this compute (inbox buffer = seq & 1)
    cudaStream_t stream = nullptr, own_stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaGraphExec_t loop_graph[2] = {nullptr, nullptr};   // per starting ping-pong parity
    int h_error = 0;
    bool use_tma = false;         // subcycle kernel v2 (TMA ring) instead of v1 (register sweep)
    bool shared_stream = false;   // inside evp_gpu_compute_group with more than one rank
    int sm_count = 148;
    CUtensorMap mapS, mapS2, mapT1, mapT3, mapT4, mapT10, mapU, mapU8, mapM;
    // device ice_strength (evp_gpu_upload_itd): aice0, aicen(ncat), vicen(ncat) planes of the arena's
    // pitch in their own allocation; strength_from_itd = the last upload had in->strength == NULL
    double *itd = nullptr;
    evp_itd_t itd_cfg;
    bool itd_ready = false, strength_from_itd = false;
    // horizontal transport, first stage (evp_gpu_set_transport_grid): dxu, dyu, HTN, HTE planes + the stop word
    double *tr = nullptr;
    int *tr_stop = nullptr;
    int h_tr_stop = 0;
    // horizontal transport, field reconstruction (evp_gpu_set_transport_tracers): hm xav yav xxav yyav, mm(0:ncat),
    // tm(ntrace, ncat), mc mx my (0:ncat), tc tx ty (ntrace, ncat) planes of the arena's pitch in one allocation
    double *rc = nullptr;
    ReconArgs rc_args;
    int *rc_stop = nullptr;       // update_fields: per category 0 or 1 + r * nxb + c of the last cell with a negative new area
    int h_rc_stop[256];
    bool rc_loaded = false, rc_remapped = false;
    // transport_remap: the ice state aice0, aicen, vicen, vsnon (ncat each), trcrn (ncat * ntrcr) as planes
    double *st = nullptr;
    bool st_loaded = false, st_done = false;
    cudaEvent_t tev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
                                                          // reconstruction: start, kernel done, ghost updates done [0..2];
                                                          // whole remap: start, end [3, 4]; departure points [5..7]
    double tr_ms[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};     // departure kernels / halos, reconstruction kernel / halos, fluxes + update, whole remap
    bool have_ht = false;         // HTE / HTN were given: compact metric rows are possible
    // EAP (kdyn = 2, evp_gpu_set_eap_tables): 6 lookup tables of kEapTableElems doubles (padded to kEapTablePitch), then
    // kEapPlanes planes of the arena's pitch: a11_1..4, a12_1..4, a11, a12, e11 e12 e22 s11 s12 s22 yieldstress11 / 12 / 22,
    // and the work planes of the two-stage subcycle (rotation of the tensor, yield stresses, alphar; evp_eap.cuh)
    double *eap = nullptr;
    bool eap_tables_set = false;
    double *eap_plane(int p) const { return eap + 6 * (size_t)kEapTablePitch + (size_t)p * plane_elems; }
    evp_stats_t stats;
    unsigned long long h_counts[2];

    double *plane(int p, int rank = -1) const
    {
        char *base = (rank < 0 || rank == cfg.rank) ? arena : peer_arena[rank];
        return reinterpret_cast<double *>(base + off_planes) + (size_t)p * plane_elems;
    }
    double *inbox(int buf, int rank = -1) const
    {
        char *base = (rank < 0 || rank == cfg.rank) ? arena : peer_arena[rank];
        return reinterpret_cast<double *>(base + off_inbox) + (size_t)buf * inbox_doubles(ib);
    }
    double *inbox_wide(int buf, int rank = -1) const
    {
        char *base = (rank < 0 || rank == cfg.rank) ? arena : peer_arena[rank];
        return reinterpret_cast<double *>(base + off_inbox_wide) + (size_t)buf * kWideSlots * inbox_doubles(ib);
    }
    unsigned long long *flags(int rank = -1) const   // [0..kMaxRanks) slots, [kMaxRanks] epoch
    {
        char *base = (rank < 0 || rank == cfg.rank) ? arena : peer_arena[rank];
        return reinterpret_cast<unsigned long long *>(base + off_flags);
    }
    int *dev_error() const { return reinterpret_cast<int *>(arena + off_flags + 1024); }
    int *iplane(int p) const   // 0 tmask, 1 umask, 2 iceumask, 3 icetmask(int)
    {
        return reinterpret_cast<int *>(arena + off_int) + (size_t)p * plane_elems;
    }
    unsigned char *mask() const { return reinterpret_cast<unsigned char *>(arena + off_mask); }
    unsigned char *tflags() const { return reinterpret_cast<unsigned char *>(arena + off_tflags); }
    unsigned char *gflags() const { return reinterpret_cast<unsigned char *>(arena + off_gflags); }
    int *queue(int ksub) const { return reinterpret_cast<int *>(arena + off_queue) + ksub; }
    unsigned long long *trace(int ksub) const
    {
        return trace_on ? reinterpret_cast<unsigned long long *>(arena + off_trace) + 8 * (size_t)ksub : nullptr;
    }
    int *done(int ksub) const { return reinterpret_cast<int *>(arena + off_queue) + cfg.ndte + ksub; }
    int *done_bnd(int ksub) const { return reinterpret_cast<int *>(arena + off_queue) + 2 * cfg.ndte + ksub; }
    int *item_count() const { return reinterpret_cast<int *>(arena + off_queue) + 3 * cfg.ndte; }   // [2]
    int *item_list() const { return reinterpret_cast<int *>(arena + off_queue) + 3 * cfg.ndte + 4; }
    unsigned long long *ll_box(int rank = -1) const
    {
        char *base = (rank < 0 || rank == cfg.rank) ? arena : peer_arena[rank];
        return reinterpret_cast<unsigned long long *>(base + off_ll);
    }
    size_t ll_bytes() const { return (size_t)kLLBufs * 2 * nyb * 4 * sizeof(unsigned long long); }
    unsigned long long *lflags(int rank = -1) const
    {
        char *base = (rank < 0 || rank == cfg.rank) ? arena : peer_arena[rank];
        return reinterpret_cast<unsigned long long *>(base + off_lflags);
    }
    unsigned long long *counts() const
    {
        return reinterpret_cast<unsigned long long *>(arena + off_counts);
    }
    int west() const { return (cfg.rank + cfg.nranks - 1) % cfg.nranks; }
    int east() const { return (cfg.rank + 1) % cfg.nranks; }
    bool west_closed() const { return cfg.ew_boundary != EVP_EW_CYCLIC && cfg.rank == 0; }
    bool east_closed() const { return cfg.ew_boundary != EVP_EW_CYCLIC && cfg.rank == cfg.nranks - 1; }
};

namespace {

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// set_evp_parameters, source/ice_dyn_shared.F90:209-257; constants
// drivers/cice/ice_constants.F90:142-148 are computed, not literals.
void set_evp_parameters(const evp_config_t &c, Consts &k, double xmin = 1.0)
{
    const double c1 = 1.0, c2 = 2.0, p5 = 0.5, eyc = 0.36;
    k.p166 = c1 / 6.0; k.p333 = c1 / 3.0; k.p111 = c1 / 9.0;
    k.p055 = k.p111 * p5; k.p027 = k.p055 * p5; k.p222 = c2 / 9.0;
    const double dte = c.dt_init / (double)c.ndte;
    const double dtei = c1 / dte;
    k.ecci = 0.25;
    const double tdamp2 = c2 * eyc * c.dt_init;
    const double dte2T = dte / tdamp2;
    if (c.revised_evp) {            // Bouillon et al. 2013 (:225-233); xmin = min(dxt, dyt) over tmask
        const double Se = 0.86, xi = 5.5e-3;
        const double gamma = 0.25 * 1.e11 * c.dt_init;
        k.revp = c1;
        k.arlx1i = c2 * xi / Se;
        k.brlx = c2 * Se * xi * gamma / (xmin * xmin);
    } else {
        k.revp = 0.0;
        k.arlx1i = dte2T;
        k.brlx = c.dt_init * dtei;
    }
    k.denom1 = c1 / (c1 + k.arlx1i);
    k.cosw = c.cosw; k.sinw = c.sinw; k.rhow = c.rhow;
}

// L2 promotion of the plane boxes: box rows start on 32-byte sectors but not on 128/256-byte lines,
// so any promotion fetches bytes of the neighbouring strips (measured: +20 % DRAM reads with 256 B)
#ifndef EVP_TMA_L2PROMO
#define EVP_TMA_L2PROMO CU_TENSOR_MAP_L2_PROMOTION_NONE
#endif
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int tensor_map_encoder(EncodeTiledFn *enc)
{
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || !fn || qres != cudaDriverEntryPointSuccess) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "cuTensorMapEncodeTiled is not available from this driver");
        return EVP_ERR_CUDA;
    }
    *enc = reinterpret_cast<EncodeTiledFn>(fn);
    return EVP_OK;
}

// Tensor maps of the subcycle kernel v2: ONE 3-D tensor (column, row, plane) over the whole plane
// arena, boxes of 128 columns x 1 row x depth planes; a 2-D byte tensor for the activity mask.
int make_tensor_maps(evp_handle *h)
{
    EncodeTiledFn enc = nullptr;
    TRY(tensor_map_encoder(&enc));
    const cuuint64_t dims[3] = {(cuuint64_t)h->ld, (cuuint64_t)h->nyb, (cuuint64_t)NPLANES};
    const cuuint64_t strides[2] = {(cuuint64_t)h->ld * 8, (cuuint64_t)h->plane_elems * 8};
    const cuuint32_t estr[3] = {1, 1, 1};
    struct { CUtensorMap *m; int depth; } maps[8] = {{&h->mapS, kStateDepth}, {&h->mapS2, 2}, {&h->mapT1, 1}, {&h->mapT3, 3},
                                                     {&h->mapT4, 4}, {&h->mapT10, 10}, {&h->mapU, kUDepth},
                                                     {&h->mapU8, kUDepth - 2}};
    for (auto &mm : maps) {
        const cuuint32_t box[3] = {(cuuint32_t)kTmaBoxW, 1, (cuuint32_t)mm.depth};
        CUresult r = enc(mm.m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, h->plane(0), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)EVP_TMA_L2PROMO,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            snprintf(g_cuda_err, sizeof(g_cuda_err), "cuTensorMapEncodeTiled(depth %d) failed: %d", mm.depth, (int)r);
            return EVP_ERR_CUDA;
        }
    }
    const cuuint64_t mdims[2] = {(cuuint64_t)h->ldm, (cuuint64_t)h->nyb};
    const cuuint64_t mstr[1] = {(cuuint64_t)h->ldm};
    const cuuint32_t mbox[2] = {(cuuint32_t)kMaskBoxW, 1};
    CUresult r = enc(&h->mapM, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, h->mask(), mdims, mstr, mbox, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "cuTensorMapEncodeTiled(mask) failed: %d", (int)r);
        return EVP_ERR_CUDA;
    }
    CK(cudaFuncSetAttribute(evp_subcycle_tma_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            kTmaStages * kStageBytes));
    CK(cudaFuncSetAttribute(evp_subcycle_tma_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            kTmaStages * kStageBytes));
    CK(cudaFuncSetAttribute(evp_subcycle_tma_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            kTmaStages * kStageBytes));
    CK(cudaFuncSetAttribute(evp_subcycle_tma_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            kTmaStages * kStageBytes));
    CK(cudaFuncSetAttribute(evp_loop_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            kTmaStages * kStageBytes));
    CK(cudaFuncSetAttribute(evp_loop_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            kTmaStages * kStageBytes));
    return EVP_OK;
}

dim3 grid2d(const evp_handle *h, int bx = 128) { return dim3((h->nxb + bx - 1) / bx, h->nyb, 1); }

}  // namespace

// =========================================================================================
extern "C" {

const char *evp_gpu_strerror(int status)
{
    switch (status) {
        case EVP_OK: return "ok";
        case EVP_ERR_ARG: return "evp_gpu: invalid argument (null pointer or inconsistent block sizes)";
        case EVP_ERR_UNSUPPORTED:
            return "evp_gpu: unsupported configuration (needs kdyn=1 (EVP) or 2 (EAP, classic relaxation), "
                   "tripole only with cyclic E-W, nranks <= 8, slabs >= 2 columns)";
        case EVP_ERR_CUDA: return "evp_gpu: CUDA runtime error (see evp_gpu_last_cuda_error)";
        case EVP_ERR_STATE: return "evp_gpu: call order violated (init -> set_grid -> [set_peers] -> run)";
        case EVP_ERR_NO_DEVICE: return "evp_gpu: no usable CUDA device; this library has no CPU fallback";
        case EVP_ERR_PEER: return "evp_gpu: multi-GPU peer mapping failed";
        case EVP_ERR_TOO_LARGE:
            return "evp_gpu: block larger than the static scratch of an allocation-free entry point "
                   "(EVP_STATIC_MAX_CELLS; see evp_gpu_last_cuda_error)";
        default: return "evp_gpu: unknown status";
    }
}

const char *evp_gpu_last_cuda_error(void) { return g_cuda_err; }

int evp_gpu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int evp_gpu_init(const evp_config_t *cfg, evp_handle_t **out)
{
    if (!cfg || !out) return EVP_ERR_ARG;
    *out = nullptr;
    if ((cfg->kdyn != 1 && cfg->kdyn != 2) || cfg->revised_evp < 0 || cfg->revised_evp > 1) return EVP_ERR_UNSUPPORTED;
    if (cfg->kdyn == 2 && cfg->revised_evp) return EVP_ERR_UNSUPPORTED;
    if (cfg->revised_evp && cfg->nranks > 1 && !(cfg->xmin > 0.0)) return EVP_ERR_ARG;   // needs the GLOBAL xmin
    if (cfg->nranks < 1 || cfg->nranks > kMaxRanks || cfg->rank < 0 || cfg->rank >= cfg->nranks)
        return EVP_ERR_UNSUPPORTED;
    if (cfg->nx_global < 2 || cfg->ny_global < 3 || cfg->ndte < 1 || !(cfg->dt_init > 0.0)) return EVP_ERR_ARG;
    // x-slabs: equal widths (slab_start all zero; nranks must divide nx_global) or the caller's
    // own boundaries (e.g. balanced by ocean cells), slab_start[r] = first global column of rank r,
    // 0-based, slab_start[0] = 0
    int start[kMaxRanks + 1];
    bool given = false;
    for (int r = 1; r < cfg->nranks; ++r) given = given || cfg->slab_start[r] != 0;
    if (given) {
        for (int r = 0; r < cfg->nranks; ++r) start[r] = cfg->slab_start[r];
        if (start[0] != 0) return EVP_ERR_ARG;
    } else {
        if (cfg->nx_global % cfg->nranks) return EVP_ERR_UNSUPPORTED;
        for (int r = 0; r < cfg->nranks; ++r) start[r] = r * (cfg->nx_global / cfg->nranks);
    }
    start[cfg->nranks] = cfg->nx_global;
    int maxw = 0, minw = cfg->nx_global;
    for (int r = 0; r < cfg->nranks; ++r) {
        const int w = start[r + 1] - start[r];
        if (w < 2) return EVP_ERR_UNSUPPORTED;
        if (w > maxw) maxw = w;
        if (w < minw) minw = w;
    }
    if (cfg->nx_block != start[cfg->rank + 1] - start[cfg->rank] + 2 || cfg->ny_block != cfg->ny_global + 2)
        return EVP_ERR_ARG;
    if (cfg->ns_boundary == EVP_NS_TRIPOLE && cfg->ew_boundary != EVP_EW_CYCLIC) return EVP_ERR_UNSUPPORTED;
    if (cfg->ns_boundary < 0 || cfg->ns_boundary > 2 || cfg->ew_boundary < 0 || cfg->ew_boundary > 1)
        return EVP_ERR_UNSUPPORTED;

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaGetDeviceCount: no CUDA device visible");
        return EVP_ERR_NO_DEVICE;
    }
    if (cfg->device < 0 || cfg->device >= ndev) return EVP_ERR_NO_DEVICE;
    CK(cudaSetDevice(cfg->device));

    evp_handle *h = new (std::nothrow) evp_handle();
    if (!h) return EVP_ERR_ARG;
    h->cfg = *cfg;
    {
        const char *kv = getenv("EVP_B200_KERNEL");          // "tma" (default) | "sweep"
        h->use_tma = !(kv && strcmp(kv, "sweep") == 0);
        const char *lv = getenv("EVP_B200_LOOP");            // "1" (default) | "0": one launch per subcycle
        h->use_loop = h->use_tma && !(lv && strcmp(lv, "0") == 0);
        // EAP: the register-sweep kernel with stress_eap as its cell function (evp_eap.cuh), one launch per subcycle
        if (cfg->kdyn == 2) h->use_tma = h->use_loop = false;
    }
    set_evp_parameters(*cfg, h->k);
    h->nxb = cfg->nx_block; h->nyb = cfg->ny_block;
    // rows start on 128-byte boundaries (TMA, sectors); the pitch is that of the widest slab so that
    // every rank's arena has the same layout (peer plane = peer base + the same offset)
    h->ld = (int)align_up((size_t)maxw + 2, 16);
    h->ldm = h->ld;
    h->plane_elems = (size_t)h->ld * h->nyb;
    Geom &g = h->g;
    g.nxl = h->nxb - 2; g.ny = h->nyb - 2; g.ld = h->ld; g.nx_global = cfg->nx_global;
    g.lo = start[cfg->rank]; g.nranks = cfg->nranks; g.rank = cfg->rank;
    for (int r = 0; r <= kMaxRanks; ++r) g.start[r] = r <= cfg->nranks ? start[r] : cfg->nx_global;
    g.tripole = cfg->ns_boundary == EVP_NS_TRIPOLE; g.ew_cyclic = cfg->ew_boundary == EVP_EW_CYCLIC;
    g.inbox_ld = h->ld;
    h->ib.inbox_ld = g.inbox_ld;
    h->ib.col_ld = (int)align_up((size_t)h->nyb, 16);

    size_t off = 0;
    h->off_planes = off; off = align_up(off + (size_t)NPLANES * h->plane_elems * sizeof(double), 256);
    h->off_inbox = off;  off = align_up(off + 2 * inbox_doubles(h->ib) * sizeof(double), 256);
    h->off_flags = off;  off = align_up(off + 4096, 256);
    h->off_int = off;    off = align_up(off + 4 * h->plane_elems * sizeof(int), 256);
    h->off_mask = off;   off = align_up(off + (size_t)h->ldm * h->nyb, 256);
    h->off_counts = off; off = align_up(off + 64, 256);
    h->tstrips = (g.nxl + kTmaUCols - 1) / kTmaUCols;
    h->tflags_ld = (int)align_up((size_t)h->nyb + 4, 16);
    {   // work items of the TMA kernels: (strip of 124 U columns) x (chunk of t_rpc rows).  The chunk height is
        // derived from the WIDEST slab so that every rank cuts its rows the same way: the persistent loop kernel
        // matches a tile with its E-W neighbour on the next rank by chunk number.
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cfg->device);
        const long long slots = (long long)kTmaBlocksPerSM * sms;
        // launch per subcycle: about 4 items per resident CTA (drawn heaviest first, so land / open-ocean items even
        // out the ragged end of a launch), at most 32 rows, at least 4.  Loop kernel: no ragged end to even out, so
        // items as tall as possible (each costs one redundant T row and one velocity-only stage) while every rank
        // keeps about 1.5 x as many items per subcycle as resident CTAs -- a tile's neighbours of the previous
        // subcycle must have left the CTAs by the time it is drawn (measured on the N = 8 per-GPU proxy, 4 strips:
        // 75.4 us at 12-16 rows, 83 at 20, 85 at 4; tp01: 520 / 511 / 507 us at 16 / 32 / 48 rows)
        const long long wstrips = (maxw + kTmaUCols - 1) / kTmaUCols, nstrips = (minw + kTmaUCols - 1) / kTmaUCols;
        long long rpc;
        if (h->use_loop) {
            rpc = (2 * nstrips * g.ny) / (3 * slots);
            rpc = rpc < 2 ? 2 : (rpc > 48 ? 48 : rpc);
            // a grid with fewer one-row items than resident CTAs (gx3: 116): one row per item.  On the L2-resident grids a
            // subcycle takes a fixed chain + about 2 us per row of an item (profiles/r02_small_grids.md); measured on one
            // B200: gx3 8.86 -> 7.15 us per subcycle, but gx1 (1152 items) 12.0 -> 14.0, so only below the slot count
            if (cfg->nranks == 1 && nstrips * g.ny <= slots) rpc = 1;
        } else {
            rpc = (wstrips * g.ny + 4 * slots - 1) / (4 * slots);
            rpc = rpc < 4 ? 4 : (rpc > 32 ? 32 : rpc);
        }
        if (const char *ev = getenv("EVP_B200_RPC")) rpc = atoi(ev);      // tuning experiments
        rpc = rpc < 1 ? 1 : (rpc > 64 ? 64 : rpc);
        h->t_rpc = (int)rpc;
        h->t_chunks = (g.ny + h->t_rpc - 1) / h->t_rpc;
        h->t_items = h->tstrips * h->t_chunks;
    }
    h->off_ll = off;     off = align_up(off + h->ll_bytes(), 256);      // (same size and offset on every rank)
    h->off_inbox_wide = off; off = align_up(off + 2 * (size_t)kWideSlots * inbox_doubles(h->ib) * sizeof(double), 256);   // (the same)
    h->off_lflags = off; off = align_up(off + (16 + 2 * (size_t)h->t_chunks + h->t_items) * sizeof(unsigned long long), 256);
    h->off_tflags = off; off = align_up(off + (size_t)h->tstrips * h->tflags_ld, 256);
    h->off_gflags = off; off = align_up(off + (size_t)h->tstrips * h->tflags_ld, 256);
    h->trace_on = getenv("EVP_B200_TRACE") != nullptr;
    h->off_trace = off;  off = align_up(off + (size_t)cfg->ndte * 8 * sizeof(unsigned long long), 256);
    h->off_queue = off;  off = align_up(off + (3 * (size_t)cfg->ndte + 4 + h->t_items) * sizeof(int), 256);
    h->arena_bytes = off;
    cudaError_t e = cudaMalloc(&h->arena, h->arena_bytes);
    if (e != cudaSuccess) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaMalloc(%zu bytes): %s", h->arena_bytes, cudaGetErrorString(e));
        h->arena = nullptr;
        evp_gpu_finalize(h);
        return EVP_ERR_CUDA;
    }
    CKH(cudaMemset(h->arena, 0, h->arena_bytes));
    CKH(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    h->stream = h->own_stream;
    for (auto &ev : h->ev) CKH(cudaEventCreate(&ev));
    memset(&h->stats, 0, sizeof(h->stats));
    {
        cudaDeviceProp prop;
        CKH(cudaGetDeviceProperties(&prop, cfg->device));
        h->sm_count = prop.multiProcessorCount;
        if (h->use_tma) { int rc = make_tensor_maps(h); if (rc != EVP_OK) { evp_gpu_finalize(h); return rc; } }
    }
    h->peer_arena[cfg->rank] = h->arena;
    h->peers_set = cfg->nranks == 1;
    *out = h;
    return EVP_OK;
}

int evp_gpu_finalize(evp_handle_t *h)
{
    if (!h) return EVP_ERR_ARG;
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    for (auto &gr : h->loop_graph) if (gr) cudaGraphExecDestroy(gr);
    for (int r = 0; r < kMaxRanks; ++r)
        if (h->peer_opened[r]) cudaIpcCloseMemHandle(h->peer_arena[r]);
    for (auto &ev : h->ev) if (ev) cudaEventDestroy(ev);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    if (h->itd) cudaFree(h->itd);
    if (h->tr) cudaFree(h->tr);
    if (h->rc) cudaFree(h->rc);
    if (h->st) cudaFree(h->st);
    if (h->eap) cudaFree(h->eap);
    for (auto &ev : h->tev) if (ev) cudaEventDestroy(ev);
    if (h->stage) cudaFree(h->stage);
    if (h->arena) cudaFree(h->arena);
    for (void *p : h->pinned) cudaHostUnregister(p);
    delete h;
    return EVP_OK;
}

int evp_gpu_pin_host_arrays(evp_handle_t *h, const evp_in_t *in, const evp_inout_t *io, const evp_out_t *out)
{
    if (!h) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    const size_t cells = (size_t)h->nxb * h->nyb;
    auto pin = [h](const void *cp, size_t bytes) -> int {
        void *p = const_cast<void *>(cp);
        if (!p) return EVP_OK;
        for (void *q : h->pinned) if (q == p) return EVP_OK;
        const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterDefault);
        if (e == cudaErrorHostMemoryAlreadyRegistered) { (void)cudaGetLastError(); return EVP_OK; }   // the caller's own
        if (e != cudaSuccess) {
            snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaHostRegister(%zu bytes): %s", bytes, cudaGetErrorString(e));
            (void)cudaGetLastError();
            return EVP_ERR_CUDA;
        }
        h->pinned.push_back(p);
        return EVP_OK;
    };
    const size_t r8 = cells * sizeof(double), i4 = cells * sizeof(int);
    if (in) {
        const double *f[14] = {in->aice, in->vice, in->vsno, in->aice_init, in->strairxT, in->strairyT, in->uocn,
                               in->vocn, in->ss_tltx, in->ss_tlty, in->Cdn_ocn, in->strength, in->strax, in->stray};
        for (const double *p : f) TRY(pin(p, r8));
    }
    if (io) {
        TRY(pin(io->uvel, r8)); TRY(pin(io->vvel, r8)); TRY(pin(io->fm, r8)); TRY(pin(io->iceumask, i4));
        for (int c = 0; c < 4; ++c) { TRY(pin(io->stressp[c], r8)); TRY(pin(io->stressm[c], r8)); TRY(pin(io->stress12[c], r8)); }
    }
    if (out) {
        double *f[19] = {out->strairx, out->strairy, out->strtltx, out->strtlty, out->strocnx, out->strocny, out->strintx,
                         out->strinty, out->strocnxT, out->strocnyT, out->divu, out->shear, out->rdg_conv, out->rdg_shear,
                         out->prs_sig, out->uvel_init, out->vvel_init, out->tmass, out->strength};
        for (double *p : f) TRY(pin(p, r8));
        TRY(pin(out->icetmask, i4));
    }
    return EVP_OK;
}

int evp_gpu_synchronize(evp_handle_t *h)
{
    if (!h) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

}  // extern "C"

// ---- host <-> device copies: dense staging + one repack kernel per batch (evp_prep.cuh) ----
static int stage_ready(evp_handle *h)
{
    if (h->stage) return EVP_OK;
    const size_t bytes = (size_t)kRepackMax * h->nxb * h->nyb * sizeof(double);
    cudaError_t e = cudaMalloc(&h->stage, bytes);
    if (e != cudaSuccess) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaMalloc(staging, %zu bytes): %s", bytes, cudaGetErrorString(e));
        return EVP_ERR_CUDA;
    }
    return EVP_OK;
}
static char *stage_slot(evp_handle *h, size_t k) { return h->stage + k * (size_t)h->nxb * h->nyb * sizeof(double); }

template <bool TO_PLANES>
static int repack(evp_handle *h)
{
    if (h->batch_dst.empty()) return EVP_OK;
    RepackArgs a;
    memset(&a, 0, sizeof(a));
    a.n = (int)h->batch_dst.size(); a.nxb = h->nxb; a.nyb = h->nyb; a.ld = h->ld;
    for (int k = 0; k < a.n; ++k) {
        a.pitched[k] = h->batch_dst[k].first; a.dense[k] = stage_slot(h, k); a.is_int[k] = (unsigned char)h->batch_dst[k].second;
    }
    dim3 grid((h->nxb + 127) / 128, h->nyb, a.n);
    auto kern = repack_kernel<TO_PLANES>;
    kern<<<grid, 128, 0, h->stream>>>(a);
    CK(cudaGetLastError());
    return EVP_OK;
}
// H2D: queue the dense copy of one array; h2d_flush() moves the batch into the planes
static int h2d_any(evp_handle *h, void *plane, const void *src, int is_int)
{
    if (!src) return EVP_ERR_ARG;
    TRY(stage_ready(h));
    if ((int)h->batch_dst.size() == kRepackMax) { TRY(repack<true>(h)); h->batch_dst.clear(); }
    const size_t bytes = (size_t)h->nxb * h->nyb * (is_int ? sizeof(int) : sizeof(double));
    CK(cudaMemcpyAsync(stage_slot(h, h->batch_dst.size()), src, bytes, cudaMemcpyHostToDevice, h->stream));
    h->batch_dst.emplace_back(plane, is_int);
    return EVP_OK;
}
static int h2d(evp_handle *h, int p, const double *src) { return h2d_any(h, h->plane(p), src, 0); }
static int h2d_i(evp_handle *h, int p, const int *src) { return h2d_any(h, h->iplane(p), src, 1); }
static int h2d_flush(evp_handle *h)
{
    const int rc = repack<true>(h);
    h->batch_dst.clear();
    return rc;
}
// D2H: collect (plane, host pointer) pairs, then one repack kernel and one dense copy per array
struct D2HBatch {
    evp_handle *h;
    std::vector<void *> host;
    int add(void *plane, void *dst, int is_int)
    {
        if (!dst) return EVP_OK;
        if ((int)h->batch_dst.size() == kRepackMax) TRY(flush());
        h->batch_dst.emplace_back(plane, is_int);
        host.push_back(dst);
        return EVP_OK;
    }
    int r8(int p, double *dst) { return add(h->plane(p), dst, 0); }
    int i4(int p, int *dst) { return add(h->iplane(p), dst, 1); }
    int flush()
    {
        if (h->batch_dst.empty()) return EVP_OK;
        TRY(stage_ready(h));
        TRY(repack<false>(h));
        for (size_t k = 0; k < host.size(); ++k) {
            const size_t bytes = (size_t)h->nxb * h->nyb * (h->batch_dst[k].second ? sizeof(int) : sizeof(double));
            CK(cudaMemcpyAsync(host[k], stage_slot(h, k), bytes, cudaMemcpyDeviceToHost, h->stream));
        }
        // the staging slots are reused by the next batch: it is enqueued behind these copies on the same stream
        h->batch_dst.clear();
        host.clear();
        return EVP_OK;
    }
};

extern "C" {

int evp_gpu_set_grid(evp_handle_t *h, const evp_grid_t *gr)
{
    if (!h || !gr) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    // a captured loop holds the constants and flags derived from the grid: drop it
    for (auto &ge : h->loop_graph) if (ge) { CK(cudaGraphExecDestroy(ge)); ge = nullptr; }
    const double *src[14] = {gr->dxt, gr->dyt, gr->dxhy, gr->dyhx, gr->cxp, gr->cyp, gr->cxm, gr->cym,
                             gr->tarea, gr->uarea, gr->tarear, gr->uarear, gr->tinyarea, gr->fcor};
    const int dstp[14] = {G_DXT, G_DYT, G_DXHY, G_DYHX, G_CXP, G_CYP, G_CXM, G_CYM,
                          G_TAREA, G_UAREA, G_TAREAR, G_UAREAR, G_TINYAREA, G_FCOR};
    for (int p = 0; p < 14; ++p) TRY(h2d(h, dstp[p], src[p]));
    if (h->cfg.revised_evp) {
        // xmin = min(global_minval(dxt, tmask), global_minval(dyt, tmask)) (ice_dyn_shared.F90:220-223):
        // given by the host when the domain is split over tasks, else taken from this block
        double xmin = h->cfg.xmin;
        if (!(xmin > 0.0)) {
            xmin = 1.0e300;
            for (int j = 1; j <= h->g.ny; ++j)
                for (int i = 1; i <= h->g.nxl; ++i) {
                    const size_t q = (size_t)j * h->nxb + i;
                    if (gr->tmask[q]) {
                        if (gr->dxt[q] < xmin) xmin = gr->dxt[q];
                        if (gr->dyt[q] < xmin) xmin = gr->dyt[q];
                    }
                }
        }
        set_evp_parameters(h->cfg, h->k, xmin);
    }
    h->have_ht = gr->HTE && gr->HTN;
    CK(cudaMemsetAsync(h->gflags(), 0, (size_t)h->tstrips * h->tflags_ld, h->stream));
    if (h->have_ht) {
        // compact metric rows: where the 8 metric terms (and tinyarea) of a row segment are
        // reproduced bit for bit from HTE / HTN (ice_grid.F90:341-367,1456,1536), the subcycle
        // kernel fetches 2 planes instead of 8 (9) and recomputes; every other segment (the
        // extrapolated row j = 1, halo-filled ghost rows, AusCOM's file tarea ...) keeps the planes
        TRY(h2d(h, G_HTE, gr->HTE));
        TRY(h2d(h, G_HTN, gr->HTN));
        TRY(h2d_flush(h));
        if (h->use_tma) {
            MetricCheckArgs m;
            m.hte = h->plane(G_HTE); m.htn = h->plane(G_HTN);
            const int mp[8] = {G_DXT, G_DYT, G_DXHY, G_DYHX, G_CXP, G_CYP, G_CXM, G_CYM};
            for (int q = 0; q < 8; ++q) m.met[q] = h->plane(mp[q]);
            m.tinyarea = h->plane(G_TINYAREA);
            m.ld = h->ld; m.nxl = h->g.nxl; m.ny = h->g.ny; m.flags = h->gflags(); m.flags_ld = h->tflags_ld;
            dim3 grid(h->tstrips, (h->g.ny + 4 + 3) / 4);
            metric_flags_kernel<<<grid, 128, 0, h->stream>>>(m);
            CK(cudaGetLastError());
        }
    }
    TRY(h2d_i(h, 0, gr->tmask));
    TRY(h2d_i(h, 1, gr->umask));
    TRY(h2d_flush(h));
    CK(cudaStreamSynchronize(h->stream));
    h->grid_set = true;
    return EVP_OK;
}

// io == NULL: only the per-call inputs go up, the dynamic state stays as the last evp left it on the device
int evp_gpu_upload(evp_handle_t *h, const evp_in_t *in, const evp_inout_t *io)
{
    if (!h || !in) return EVP_ERR_ARG;
    if (!h->grid_set) return EVP_ERR_STATE;
    if (!io && !h->uploaded) return EVP_ERR_STATE;               // no resident state yet
    CK(cudaSetDevice(h->cfg.device));
    const double *src[12] = {in->aice, in->vice, in->vsno, in->aice_init, in->strairxT, in->strairyT,
                             in->uocn, in->vocn, in->ss_tltx, in->ss_tlty, in->Cdn_ocn, in->strength};
    const int dstp[12] = {I_AICE, I_VICE, I_VSNO, I_AICE_INIT, I_STRAIRXT, I_STRAIRYT,
                          I_UOCN, I_VOCN, I_SSTLTX, I_SSTLTY, I_CW, I_STRENGTH};
    if (!in->strength && !h->itd_ready) return EVP_ERR_ARG;      // neither the host's strength nor an ITD
    if ((in->strax == nullptr) != (in->stray == nullptr)) return EVP_ERR_ARG;
    h->strength_from_itd = in->strength == nullptr;
    // ACCESS, .not. calc_strair (ice_dyn_evp.F90:220-229): strairxT/yT reach no output of the call then
    // (evp_prep1's copy of them is overwritten), so the coupler's U-grid stress takes their planes
    h->strair_given = in->strax != nullptr;
    if (h->strair_given) { src[4] = in->strax; src[5] = in->stray; }
    for (int p = 0; p < 12; ++p) {
        if (dstp[p] == I_STRENGTH && h->strength_from_itd) continue;      // computed on the device
        // the sea-surface slope is read only where the tilt comes from it (#ifdef coupled,
        // ice_dyn_shared.F90:601-602): two planes less to upload otherwise; the pointers may be NULL then
        if ((dstp[p] == I_SSTLTX || dstp[p] == I_SSTLTY) && h->cfg.tilt_mode == 0) continue;
        TRY(h2d(h, dstp[p], src[p]));
    }
    if (io) {
        // dynamic state always lands in copy A
        TRY(h2d(h, S_UA, io->uvel));
        TRY(h2d(h, S_VA, io->vvel));
        for (int c = 0; c < 4; ++c) {
            TRY(h2d(h, S_STRA + c, io->stressp[c]));
            TRY(h2d(h, S_STRA + 4 + c, io->stressm[c]));
            TRY(h2d(h, S_STRA + 8 + c, io->stress12[c]));
        }
        TRY(h2d_i(h, 2, io->iceumask));
        h->parity = 0;
        h->uploaded = true;
        h->state_valid = true;
    }
    TRY(h2d_flush(h));
    return EVP_OK;
}

int evp_gpu_upload_itd(evp_handle_t *h, const evp_itd_t *itd)
{
    if (!h || !itd || !itd->aicen || !itd->vicen || !itd->aice0) return EVP_ERR_ARG;
    if (itd->ncat < 1 || itd->ncat > 16) return EVP_ERR_UNSUPPORTED;
    if (itd->kstrength < 0 || itd->kstrength > 1 || itd->krdg_partic < 0 || itd->krdg_partic > 1 ||
        itd->krdg_redist < 0 || itd->krdg_redist > 1)
        return EVP_ERR_UNSUPPORTED;
    CK(cudaSetDevice(h->cfg.device));
    const size_t nplanes = 2 * (size_t)itd->ncat + 1;
    if (!h->itd || h->itd_cfg.ncat != itd->ncat) {
        if (h->itd) { CK(cudaFree(h->itd)); h->itd = nullptr; }
        CK(cudaMalloc(&h->itd, nplanes * h->plane_elems * sizeof(double)));
    }
    h->itd_cfg = *itd;
    const size_t hostplane = (size_t)h->nxb * h->nyb;
    for (size_t p = 0; p < nplanes; ++p) {      // [0] aice0, [1 .. ncat] aicen, [ncat+1 .. 2 ncat] vicen
        const double *src = p == 0 ? itd->aice0
                          : p <= (size_t)itd->ncat ? itd->aicen + (p - 1) * hostplane
                                                   : itd->vicen + (p - 1 - itd->ncat) * hostplane;
        TRY(h2d_any(h, h->itd + p * h->plane_elems, src, 0));
    }
    TRY(h2d_flush(h));
    h->itd_ready = true;
    return EVP_OK;
}

static int download_groups(evp_handle *h, evp_inout_t *io, evp_out_t *out, unsigned mask)
{
    if (!h->uploaded) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    D2HBatch B{h, {}};
    h->batch_dst.clear();
    if (io) {
        const int U = h->parity ? S_UB : S_UA, V = h->parity ? S_VB : S_VA;
        const int S = h->parity ? S_STRB : S_STRA;
        if (mask & EVP_DL_VELOCITY) { TRY(B.r8(U, io->uvel)); TRY(B.r8(V, io->vvel)); }
        if (mask & EVP_DL_STRESS)
            for (int c = 0; c < 4; ++c) {
                TRY(B.r8(S + c, io->stressp[c]));
                TRY(B.r8(S + 4 + c, io->stressm[c]));
                TRY(B.r8(S + 8 + c, io->stress12[c]));
            }
        if (mask & EVP_DL_HISTORY) { TRY(B.r8(S_FM, io->fm)); TRY(B.i4(2, io->iceumask)); }
    }
    if (out) {
        double *dst[17] = {out->strairx, out->strairy, out->strtltx, out->strtlty, out->strocnx, out->strocny,
                           out->strintx, out->strinty, out->strocnxT, out->strocnyT, out->divu, out->shear,
                           out->rdg_conv, out->rdg_shear, out->prs_sig, out->uvel_init, out->vvel_init};
        // groups of the 17 output planes (evp_out_t order)
        static const unsigned grp[17] = {EVP_DL_HISTORY, EVP_DL_HISTORY, EVP_DL_HISTORY, EVP_DL_HISTORY, EVP_DL_HISTORY,
                                         EVP_DL_HISTORY, EVP_DL_HISTORY, EVP_DL_HISTORY, EVP_DL_OCN_T, EVP_DL_OCN_T,
                                         EVP_DL_DEFORM, EVP_DL_DEFORM, EVP_DL_DEFORM, EVP_DL_DEFORM, EVP_DL_HISTORY,
                                         EVP_DL_HISTORY, EVP_DL_HISTORY};
        for (int p = 0; p < 17; ++p)
            if (mask & grp[p]) TRY(B.r8(O_STRAIRX + p, dst[p]));
        if (mask & EVP_DL_HISTORY) {
            TRY(B.r8(W_TMASS, out->tmass));
            TRY(B.r8(W_STRENGTH, out->strength));
            TRY(B.i4(3, out->icetmask));
        }
    }
    TRY(B.flush());
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

int evp_gpu_download(evp_handle_t *h, evp_inout_t *io, evp_out_t *out)
{
    if (!h) return EVP_ERR_ARG;
    return download_groups(h, io, out, EVP_DL_ALL);
}

int evp_gpu_set_resident(evp_handle_t *h, int resident, unsigned download_mask)
{
    if (!h || (download_mask & ~(unsigned)EVP_DL_ALL)) return EVP_ERR_ARG;
    h->resident = resident != 0;
    h->dl_mask = download_mask;
    return EVP_OK;
}

int evp_gpu_invalidate_state(evp_handle_t *h)
{
    if (!h) return EVP_ERR_ARG;
    h->state_valid = false;
    return EVP_OK;
}

int evp_gpu_get_stats(evp_handle_t *h, evp_stats_t *st)
{
    if (!h || !st) return EVP_ERR_ARG;
    *st = h->stats;
    return EVP_OK;
}

}  // extern "C"

// =========================================================================================
namespace {

struct Launcher {
    evp_handle *h;
    long long n = 0;
    cudaStream_t s() const { return h->stream; }
};

// the persistent loop kernel needs its own stream per rank (a kernel must never wait for a kernel queued
// behind it): evp_gpu_compute_group, which drives all ranks through one stream, keeps a launch per subcycle
bool loop_mode(const evp_handle *h) { return h->use_loop && h->use_tma && !h->shared_stream; }

// device-side waits give up after this long (EVP_B200_SYNC_TIMEOUT_S, default 10 s): the error word is set,
// every later wait of the call returns at once and the call ends with EVP_ERR_PEER
unsigned long long sync_timeout_ns()
{
    double s = 10.0;
    if (const char *ev = getenv("EVP_B200_SYNC_TIMEOUT_S")) { const double v = atof(ev); if (v > 0.0) s = v; }
    return (unsigned long long)(s * 1.0e9);
}

void fill_sync_args(const evp_handle *h, SyncArgs &a)
{
    memset(&a, 0, sizeof(a));
    a.rank = h->cfg.rank; a.nranks = h->cfg.nranks;
    a.epoch = h->flags() + kMaxRanks;
    for (int r = 0; r < h->cfg.nranks; ++r) a.flags[r] = h->flags(r);
    a.error = h->dev_error();
    a.timeout_ns = sync_timeout_ns();
}

int launch_signal(Launcher &L)
{
    SyncArgs a;
    fill_sync_args(L.h, a);
    halo_signal_kernel<<<1, 32, 0, L.s()>>>(a);
    ++L.n;
    return EVP_OK;
}
int launch_wait(Launcher &L)
{
    SyncArgs a;
    fill_sync_args(L.h, a);
    halo_wait_kernel<<<1, 32, 0, L.s()>>>(a);
    ++L.n;
    return EVP_OK;
}

// ---- program builder: the evp(dt) driver (source/ice_dyn_evp.F90:68-510) as a list of steps ----
struct Builder {
    evp_handle *h;
    std::vector<Op> &prog;
    int seq = 0;                               // halo operations so far -> inbox buffer
    void op(std::function<int(Launcher &)> f) { prog.push_back(Op{false, std::move(f)}); }
    void sync() { if (h->cfg.nranks > 1) prog.push_back(Op{true, nullptr}); }

    // ice_HaloUpdate of one device plane: push edges / top rows into the inboxes, sync, apply
    // first_pair: kFirstPair3D for the levels of fld2 (ice_HaloUpdate3DR8), see fold_necorner_column
    void halo(int plane, int loc, double sgn, double fill, int first_pair = kFirstPair2D)
    {
        halo_p(h->plane(plane), loc, sgn, fill, first_pair);
    }
    void halo_p(double *plane, int loc, double sgn, double fill, int first_pair = kFirstPair2D)
    {
        evp_handle *hh = h;
        const int buf = (seq++) & 1;
        op([hh, plane, loc, buf](Launcher &L) {
            const Geom &g = hh->g;
            PushArgs pa;
            memset(&pa, 0, sizeof(pa));
            pa.g = g; pa.ib = hh->ib; pa.plane = plane;
            for (int r = 0; r < g.nranks; ++r) pa.inbox[r] = hh->inbox(buf, r);
            pa.west = hh->west_closed() ? -1 : hh->west();
            pa.east = hh->east_closed() ? -1 : hh->east();
            pa.ioff = loc; pa.ew = 1;
            const int nt = g.ny + (g.tripole ? g.nxl : 0);
            halo_push_kernel<<<(nt + 255) / 256, 256, 0, L.s()>>>(pa);
            ++L.n;
            return EVP_OK;
        });
        sync();
        op([hh, plane, loc, sgn, fill, buf, first_pair](Launcher &L) {
            const Geom &g = hh->g;
            const int na = g.nxl + 2 + g.ny;
            halo_apply_kernel<<<(na + 255) / 256, 256, 0, L.s()>>>(plane, g, hh->ib, hh->inbox(buf), loc,
                                                                sgn, fill, hh->west_closed(), hh->east_closed(),
                                                                first_pair);
            ++L.n;
            CK(cudaGetLastError());
            return EVP_OK;
        });
    }

    // ice_HaloUpdate (center) of `nplanes` planes `stride` elements apart, 2 * kWideSlots planes per operation
    int wide_seq = 0;
    void halo_multi(double *plane0, int nplanes, size_t stride, NegRanges neg, double fill)
    {
        evp_handle *hh = h;
        for (int first = 0; first < nplanes; first += 2 * kWideSlots) {
            const int np = nplanes - first < 2 * kWideSlots ? nplanes - first : 2 * kWideSlots;
            double *plane = plane0 + (size_t)first * stride;
            const int buf = (wide_seq++) & 1;
            op([hh, plane, np, stride, buf](Launcher &L) {
                const Geom &g = hh->g;
                PushArgs pa;
                memset(&pa, 0, sizeof(pa));
                pa.g = g; pa.ib = hh->ib; pa.plane = plane;
                for (int r = 0; r < g.nranks; ++r) pa.inbox[r] = hh->inbox_wide(buf, r);
                pa.west = hh->west_closed() ? -1 : hh->west();
                pa.east = hh->east_closed() ? -1 : hh->east();
                pa.ioff = 0; pa.ew = 1;
                const int nt = g.ny + (g.tripole ? g.nxl : 0);
                halo_push_multi_kernel<<<dim3((nt + 255) / 256, np), 256, 0, L.s()>>>(pa, stride);
                ++L.n;
                return EVP_OK;
            });
            sync();
            op([hh, plane, np, stride, first, neg, fill, buf](Launcher &L) {
                const Geom &g = hh->g;
                const int na = g.nxl + 2 + g.ny;
                halo_apply_multi_kernel<<<dim3((na + 255) / 256, np), 256, 0, L.s()>>>(
                    plane, stride, g, hh->ib, hh->inbox_wide(buf), first, neg, fill, hh->west_closed(), hh->east_closed());
                ++L.n;
                CK(cudaGetLastError());
                return EVP_OK;
            });
        }
    }

    // ice_HaloUpdate_stress(a1, a2): a2's top physical row mirrored with center offsets
    void halo_stress(int a1, int a2)
    {
        evp_handle *hh = h;
        const int buf = (seq++) & 1;
        if (hh->g.tripole) {
            op([hh, a2, buf](Launcher &L) {
                const Geom &g = hh->g;
                PushArgs pa;
                memset(&pa, 0, sizeof(pa));
                pa.g = g; pa.ib = hh->ib; pa.plane = hh->plane(a2);
                for (int r = 0; r < g.nranks; ++r) pa.inbox[r] = hh->inbox(buf, r);
                pa.west = -1; pa.east = -1;            // no ordinary neighbour copies (:3346-3353)
                pa.ioff = 0; pa.ew = 0;
                const int nt = g.ny + g.nxl;
                halo_push_kernel<<<(nt + 255) / 256, 256, 0, L.s()>>>(pa);
                ++L.n;
                return EVP_OK;
            });
            sync();
        }
        op([hh, a1, buf](Launcher &L) {
            const Geom &g = hh->g;
            const int na = g.nxl + 2 + g.ny + 2;
            halo_stress_apply_kernel<<<(na + 255) / 256, 256, 0, L.s()>>>(hh->plane(a1), g, hh->ib, hh->inbox(buf),
                                                                       hh->west_closed(), hh->east_closed());
            ++L.n;
            CK(cudaGetLastError());
            return EVP_OK;
        });
    }
};

void fill_loop_args(const evp_handle *h, int from_parity, int inbox_buf, LoopArgs &a)
{
    const Geom &g = h->g;
    memset(&a, 0, sizeof(a));
    a.nxl = g.nxl; a.ny = g.ny; a.ld = g.ld; a.ldm = h->ldm;
    {   // rows marched by one warp: long chunks amortise the redundant north row on big grids,
        // short chunks give the small (L2-resident, latency-bound) grids enough warps
        const long long strips = (g.nxl + kUColsPerWarp - 1) / kUColsPerWarp;
        const long long want_warps = 148LL * 16 * 2;
        long long rpc = (strips * g.ny + want_warps - 1) / want_warps;
        a.rows_per_chunk = (int)(rpc < 2 ? 2 : (rpc > 32 ? 32 : rpc));
    }
    a.nx_global = g.nx_global; a.lo = g.lo; a.nranks = g.nranks; a.rank = g.rank; a.tripole = g.tripole;
    for (int r = 0; r <= kMaxRanks; ++r) a.start[r] = g.start[r];
    a.nxl_west = g.start[h->west() + 1] - g.start[h->west()];
    a.inbox_ld = g.inbox_ld;
    a.mask = h->mask();
    const int U0 = from_parity ? S_UB : S_UA, V0 = from_parity ? S_VB : S_VA;
    const int U1 = from_parity ? S_UA : S_UB, V1 = from_parity ? S_VA : S_VB;
    const int T0 = from_parity ? S_STRB : S_STRA, T1 = from_parity ? S_STRA : S_STRB;
    a.u0 = h->plane(U0); a.v0 = h->plane(V0); a.u1 = h->plane(U1); a.v1 = h->plane(V1);
    for (int q = 0; q < 12; ++q) { a.s0[q] = h->plane(T0 + q); a.s1[q] = h->plane(T1 + q); }
    a.dxt = h->plane(G_DXT); a.dyt = h->plane(G_DYT); a.dxhy = h->plane(G_DXHY); a.dyhx = h->plane(G_DYHX);
    a.cxp = h->plane(G_CXP); a.cyp = h->plane(G_CYP); a.cxm = h->plane(G_CXM); a.cym = h->plane(G_CYM);
    a.tinyarea = h->plane(G_TINYAREA); a.strength = h->plane(W_STRENGTH); a.tarear = h->plane(G_TAREAR);
    a.aiu = h->plane(W_AIU); a.Cw = h->plane(I_CW); a.uocn = h->plane(I_UOCN); a.vocn = h->plane(I_VOCN);
    a.waterx = h->plane(W_WATERX); a.watery = h->plane(W_WATERY); a.forcex = h->plane(W_FORCEX);
    a.forcey = h->plane(W_FORCEY); a.umassdti = h->plane(W_UMASSDTI); a.fm = h->plane(S_FM);
    a.uarear = h->plane(G_UAREAR);
    a.divu = h->plane(O_DIVU); a.shear = h->plane(O_SHEAR); a.rdg_conv = h->plane(O_RDGCONV);
    a.rdg_shear = h->plane(O_RDGSHEAR); a.prs_sig = h->plane(O_PRSSIG);
    a.strintx = h->plane(O_STRINTX); a.strinty = h->plane(O_STRINTY);
    a.uvel_init = h->plane(O_UVELINIT); a.vvel_init = h->plane(O_VVELINIT);
    // E-W ghosts of the NEW velocity planes are written in place (self, or peer memory over
    // NVLink): the ping-pong separates these stores from every read of the same plane by one sync
    if (!h->west_closed()) { a.uW = h->plane(U1, h->west()); a.vW = h->plane(V1, h->west()); }
    if (!h->east_closed()) { a.uE = h->plane(U1, h->east()); a.vE = h->plane(V1, h->east()); }
    for (int r = 0; r < g.nranks; ++r) a.inbox[r] = h->inbox(inbox_buf, r);
    a.k = h->k;
}

void fill_eap_args(const evp_handle *h, EapArgs &e)
{
    memset(&e, 0, sizeof(e));
    for (int t = 0; t < 6; ++t) e.tab[t] = h->eap + (size_t)t * kEapTablePitch;
    for (int q = 0; q < 4; ++q) { e.a11c[q] = h->eap_plane(q); e.a12c[q] = h->eap_plane(4 + q); }
    e.a11 = h->eap_plane(8); e.a12 = h->eap_plane(9);
    for (int q = 0; q < 9; ++q) e.out[q] = h->eap_plane(10 + q);
    for (int q = 0; q < 16; ++q) e.rot[q] = h->eap_plane(19 + q);
    for (int q = 0; q < 12; ++q) e.yld[q] = h->eap_plane(35 + q);
    for (int q = 0; q < 4; ++q) e.alphar[q] = h->eap_plane(47 + q);
    // dtei = c1 / dte, dte = dt / ndte with the dt of init_evp (ice_dyn_shared.F90:209-210)
    e.dtei = 1.0 / (h->cfg.dt_init / (double)h->cfg.ndte);
}

// all ndte subcycles as one persistent kernel (evp_loop_tma.cuh): work items ordered by (subcycle, item),
// per-tile dependency counters instead of launch boundaries and cross-rank barriers
void add_loop(Builder &B, int from_parity, int nsteps)
{
    evp_handle *h = B.h;
    const int buf0 = B.seq & 1;
    B.seq += nsteps;                             // one inbox buffer per subcycle, alternating
    B.op([h, from_parity, buf0, nsteps](Launcher &L) {
        LoopTmaArgs t;
        memset(&t, 0, sizeof(t));
        fill_loop_args(h, from_parity, buf0, t.a);
        t.a.rows_per_chunk = h->t_rpc;
        const int nr = h->cfg.nranks;
        t.plane_state[0] = S_UA; t.plane_state[1] = S_UB;
        t.plane_t = kTGroup; t.plane_u = kUGroup;
        t.state[0] = h->plane(S_UA); t.state[1] = h->plane(S_UB);
        if (!h->west_closed()) { t.west_state[0] = h->plane(S_UA, h->west()); t.west_state[1] = h->plane(S_UB, h->west()); }
        if (!h->east_closed()) { t.east_state[0] = h->plane(S_UA, h->east()); t.east_state[1] = h->plane(S_UB, h->east()); }
        t.plane_elems = h->plane_elems;
        for (int b = 0; b < 2; ++b)
            for (int r = 0; r < nr; ++r) t.inbox[b][r] = h->inbox(b, r);
        t.parity0 = from_parity; t.inbox0 = buf0;
        t.strips = h->tstrips; t.chunks = h->t_chunks; t.nsteps = nsteps;
        t.has_fold = h->g.tripole ? 1 : 0;
        t.queue = h->queue(0); t.item_list = h->item_list(); t.item_count = h->item_count();
        t.tile_flags = h->tflags(); t.flags_ld = h->tflags_ld;
        t.water_is_ocn = (h->cfg.cosw == 1.0 && h->cfg.sinw == 0.0) ? 1 : 0;
        t.puny = 1.0e-11;
        unsigned long long *lf = h->lflags();
        t.topflag = lf; t.fold_flag = lf + 8; t.top_count = lf + 9;
        t.tflag = lf + 16 + 2 * (size_t)h->t_chunks;
        t.ew_self = (nr == 1 && h->g.ew_cyclic) ? 1 : 0;
        if (nr > 1 && !h->west_closed()) { t.wflag = lf + 16; t.pub_w = h->lflags(h->west()) + 16 + h->t_chunks; }
        if (nr > 1 && !h->east_closed()) { t.eflag = lf + 16 + h->t_chunks; t.pub_e = h->lflags(h->east()) + 16; }
        for (int r = 0; r < nr; ++r) t.pub_top[r] = h->lflags(r);
        t.ll_on = nr > 1 ? 1 : 0; t.nyb = h->nyb;
        t.ll_my = h->ll_box();
        if (nr > 1 && !h->west_closed()) t.ll_w = h->ll_box(h->west());
        if (nr > 1 && !h->east_closed()) t.ll_e = h->ll_box(h->east());
        t.rank = h->cfg.rank; t.nranks = nr;
        t.error = h->dev_error();
        t.timeout_ns = sync_timeout_ns();
        const CUtensorMap &mapU = t.water_is_ocn ? h->mapU8 : h->mapU;
        const int slots = kTmaBlocksPerSM * h->sm_count;
        const long long want = 2LL * (h->t_items + 1);        // small grids: two subcycles' worth of CTAs in flight
        const int grid = (int)(want < slots ? want : slots);
        const size_t smem = (size_t)kTmaStages * kStageBytes;
        auto kern = h->cfg.revised_evp ? evp_loop_tma_kernel<true> : evp_loop_tma_kernel<false>;
        kern<<<grid, (kTmaConsumers + 1) * 32, smem, L.s()>>>(h->mapS, h->mapS2, h->mapT1, h->mapT3, h->mapT4, h->mapT10,
                                                               mapU, h->mapM, t);
        ++L.n;
        return EVP_OK;
    });
    B.sync();      // every rank has finished its loop kernel: mailboxes, inboxes and folds are complete
    if (h->cfg.nranks > 1) {
        // the ghost columns of the final planes out of the mailbox rows of the loop's last subcycle
        B.op([h, from_parity, nsteps](Launcher &L) {
            const int par_end = from_parity ^ (nsteps & 1);
            const int jhi = h->g.tripole ? h->g.ny - 1 : h->g.ny;
            dim3 grid((jhi + 127) / 128, 2);
            ll_drain_kernel<<<grid, 128, 0, L.s()>>>(h->plane(par_end ? S_UB : S_UA), h->plane_elems, h->ld, h->g.nxl, h->nyb,
                                                     h->ll_box(), nsteps - 1, jhi, h->west_closed() ? 0 : 1,
                                                     h->east_closed() ? 0 : 1);
            ++L.n;
            CK(cudaGetLastError());
            return EVP_OK;
        });
    }
}

// one subcycle: stress + stepu + halo pushes fused; [sync]; then the tripole fold of the new u, v
void add_subcycle(Builder &B, int from_parity, bool last, int ksub)
{
    evp_handle *h = B.h;
    const int buf = (B.seq++) & 1;
    // N > 1 on a tripole grid with the TMA kernel: the cross-rank barrier between the subcycle
    // kernel and the fold is fused into both -- the last CTA to finish signals the peers, the fold
    // kernel waits for theirs -- 2 launches per subcycle instead of 4
    // (the fused signal relies on the boundary items coming first in the work list: not in the loop kernel's order)
    const bool fused_sync = h->cfg.nranks > 1 && h->use_tma && h->g.tripole && !loop_mode(h);
    B.op([h, from_parity, last, buf, ksub, fused_sync](Launcher &L) {
        const Geom &g = h->g;
        LoopArgs a;
        fill_loop_args(h, from_parity, buf, a);
        if (h->use_tma) {
            // v2: persistent CTAs (2 per SM) fed by a TMA ring; work item = (strip of 124 U columns,
            // chunk of rows).  Chunk height: fewest rounds x (rows + pipeline fill) over the CTAs.
            TmaArgs t;
            t.a = a;
            t.plane_state = from_parity ? S_UB : S_UA;
            t.plane_t = kTGroup; t.plane_u = kUGroup;
            t.puny = 1.0e-11;
            // waterx = uocn*cosw - vocn*sinw*sign(fm) is uocn itself (up to the sign of zero) when the
            // turning angle is zero: skip those two planes (ice_dyn_shared.F90:590-591)
            t.water_is_ocn = (h->cfg.cosw == 1.0 && h->cfg.sinw == 0.0) ? 1 : 0;
            const CUtensorMap &mapU = t.water_is_ocn ? h->mapU8 : h->mapU;
            t.strips = h->tstrips;
            t.tile_flags = h->tflags(); t.flags_ld = h->tflags_ld;
            const int slots = kTmaBlocksPerSM * h->sm_count;
            t.a.rows_per_chunk = h->t_rpc;
            t.chunks = h->t_chunks;
            t.items = h->t_items;
            t.queue = h->queue(ksub - 1); t.item_list = h->item_list(); t.item_count = h->item_count();
            t.fused_signal = fused_sync ? 1 : 0;
            t.done = h->done(ksub - 1); t.done_bnd = h->done_bnd(ksub - 1);
            fill_sync_args(h, t.sync);
            t.trace = h->trace(ksub - 1);
            const int grid = t.items < slots ? t.items : slots;
            const size_t smem = (size_t)kTmaStages * kStageBytes;
            auto kern = h->cfg.revised_evp ? (last ? evp_subcycle_tma_kernel<true, true> : evp_subcycle_tma_kernel<false, true>)
                                           : (last ? evp_subcycle_tma_kernel<true, false> : evp_subcycle_tma_kernel<false, false>);
            kern<<<grid, (kTmaConsumers + 1) * 32, smem, L.s()>>>(h->mapS, h->mapS2, h->mapT1, h->mapT3, h->mapT4,
                                                                   h->mapT10, mapU, h->mapM, t);
            ++L.n;
            return EVP_OK;
        }
        const int strips = (g.nxl + kUColsPerWarp - 1) / kUColsPerWarp;
        dim3 grid((strips + kWarpsPerBlock - 1) / kWarpsPerBlock, (g.ny + a.rows_per_chunk - 1) / a.rows_per_chunk);
        if (h->cfg.kdyn == 2) {
            // eap (source/ice_dyn_eap.F90:340-421) in two stages: the yield stresses of stress_eap, one thread per (cell,
            // corner); then relaxation + combinations + stepu in the row sweep; stepa on every 10th subcycle (:408) from the
            // stresses this subcycle wrote
            EapArgs e;
            fill_eap_args(h, e);
            dim3 g2((g.nxl + 1 + 31) / 32, g.ny + 1);
            auto kern0 = last ? eap_yield_kernel<true> : eap_yield_kernel<false>;
            kern0<<<g2, 128, 0, L.s()>>>(a, e);
            ++L.n;
            auto kern2 = last ? eap_subcycle_kernel<true> : eap_subcycle_kernel<false>;
            kern2<<<grid, kWarpsPerBlock * 32, 0, L.s()>>>(a, e);
            ++L.n;
            if (ksub % 10 == 1) {
                eap_stepa_kernel<<<g2, 128, 0, L.s()>>>(e, g.nxl, g.ny, g.ld, h->ldm, h->mask(),
                                                        h->plane(from_parity ? S_STRA : S_STRB), h->plane_elems);
                ++L.n;
            }
            CK(cudaGetLastError());
            return EVP_OK;
        }
        auto kern1 = h->cfg.revised_evp ? (last ? evp_subcycle_kernel<true, true> : evp_subcycle_kernel<false, true>)
                                        : (last ? evp_subcycle_kernel<true, false> : evp_subcycle_kernel<false, false>);
        kern1<<<grid, kWarpsPerBlock * 32, 0, L.s()>>>(a);
        ++L.n;
        return EVP_OK;
    });
    if (!fused_sync) B.sync();
    if (h->g.tripole) {
        B.op([h, from_parity, buf, fused_sync, ksub](Launcher &L) {
            const Geom &g = h->g;
            double *u1 = h->plane(from_parity ? S_UA : S_UB), *v1 = h->plane(from_parity ? S_VA : S_VB);
            SyncArgs sa;
            fill_sync_args(h, sa);
            fold_uv_kernel<<<(g.nxl + 2 + 127) / 128, 128, 0, L.s()>>>(u1, v1, g.ld, g.nxl, g.ny, g.nx_global, g.lo,
                                                                     h->inbox(buf), g.inbox_ld, sa, fused_sync ? 1 : 0,
                                                                     h->trace(ksub - 1));
            ++L.n;
            return EVP_OK;
        });
    }
}

struct Program {
    std::vector<Op> ops;
    size_t loop_begin = 0, loop_end = 0;       // [begin, end) = the ndte subcycles
    int end_parity = 0;
};

void build_program(evp_handle *h, double dt, Program &P)
{
    const evp_config_t &cfg = h->cfg;
    Builder B{h, P.ops};
    const size_t pb = h->plane_elems * sizeof(double);
    const int cur = h->parity;
    const int Ucur = cur ? S_UB : S_UA, Vcur = cur ? S_VB : S_VA, Uoth = cur ? S_UA : S_UB;

    B.sync();      // every rank has left its previous call: inbox buffers and planes are free
    B.op([h, pb](Launcher &L) {
        cudaStream_t s = L.s();
        CK(cudaEventRecord(h->ev[0], s));
        // init_history_dyn contract (ice_flux.F90:639-673) + evp's own zeroing (ice_dyn_evp.F90:174-182)
        CK(cudaMemsetAsync(h->plane(O_STRAIRX), 0, pb * 17, s));   // all output planes
        CK(cudaMemsetAsync(h->plane(S_FM), 0, pb, s));
        CK(cudaMemsetAsync(h->counts(), 0, 16, s));
        // eap zeroes its strain / stress / yield diagnostics too (source/ice_dyn_eap.F90:166-182)
        if (h->cfg.kdyn == 2) CK(cudaMemsetAsync(h->eap_plane(10), 0, pb * 9, s));
        // ghost mailbox of the loop kernel: its flags count subcycles from 1 in every call (the neighbours write it only
        // inside their loop kernels, several cross-rank barriers after this point)
        if (h->cfg.nranks > 1 && loop_mode(h)) CK(cudaMemsetAsync(h->ll_box(), 0, h->ll_bytes(), s));
        // evp_prep1 (:194)
        Prep1Args a;
        a.g = h->g; a.rhoi = h->cfg.rhoi; a.rhos = h->cfg.rhos; a.a_min = h->cfg.a_min; a.m_min = h->cfg.m_min;
        a.aice = h->plane(I_AICE); a.vice = h->plane(I_VICE); a.vsno = h->plane(I_VSNO);
        a.strairxT = h->plane(I_STRAIRXT); a.strairyT = h->plane(I_STRAIRYT); a.tmask = h->iplane(0);
        a.tmass = h->plane(W_TMASS); a.work1 = h->plane(W_1); a.work2 = h->plane(W_2);
        a.icetmask = h->plane(W_ICETMASK);
        prep1_kernel<<<grid2d(h), 128, 0, s>>>(a);
        ++L.n;
        return EVP_OK;
    });
    B.halo(W_ICETMASK, 0, +1.0, 0.0);                          // :210
    B.halo(W_1, 0, -1.0, 0.0);                                 // t2ugrid_vector (ice_grid.F90:1817)
    B.halo(W_2, 0, -1.0, 0.0);
    B.op([h, dt, cur](Launcher &L) {
        cudaStream_t s = L.s();
        const evp_config_t &cfg = h->cfg;
        {   // to_ugrid: umass, aiu, strairx, strairy (:218-241)
            ToUArgs a;
            a.g = h->g; a.tarea = h->plane(G_TAREA); a.uarea = h->plane(G_UAREA);
            a.in[0] = h->plane(W_TMASS); a.out[0] = h->plane(W_UMASS);
            a.in[1] = h->plane(I_AICE_INIT); a.out[1] = h->plane(W_AIU);
            a.in[2] = h->plane(W_1); a.out[2] = h->plane(O_STRAIRX);
            a.in[3] = h->plane(W_2); a.out[3] = h->plane(O_STRAIRY);
            to_ugrid4_kernel<<<grid2d(h), 128, 0, s>>>(a);
            ++L.n;
        }
        if (h->strair_given) {          // strairx(:,:,:) = strax(:,:,:) instead of t2ugrid_vector (:227-228)
            const size_t pbytes = h->plane_elems * sizeof(double);
            CK(cudaMemcpyAsync(h->plane(O_STRAIRX), h->plane(I_STRAIRXT), pbytes, cudaMemcpyDeviceToDevice, s));
            CK(cudaMemcpyAsync(h->plane(O_STRAIRY), h->plane(I_STRAIRYT), pbytes, cudaMemcpyDeviceToDevice, s));
        }
        if (h->strength_from_itd) {     // ice_strength on the device (:291-301), into the input plane
            const evp_itd_t &it = h->itd_cfg;
            StrengthArgs sa;
            sa.nxb = h->nxb; sa.nyb = h->nyb; sa.ld = h->ld;
            sa.ncat = it.ncat; sa.kstrength = it.kstrength; sa.krdg_partic = it.krdg_partic;
            sa.krdg_redist = it.krdg_redist; sa.mu_rdg = it.mu_rdg; sa.Cf = it.Cf;
            sa.Cp = 0.5 * cfg.gravit * (cfg.rhow - cfg.rhoi) * cfg.rhoi / cfg.rhow;
            sa.cat_stride = h->plane_elems;
            sa.aice = h->plane(I_AICE); sa.vice = h->plane(I_VICE);
            sa.aice0 = h->itd; sa.aicen = h->itd + h->plane_elems;
            sa.vicen = h->itd + (1 + (size_t)it.ncat) * h->plane_elems;
            sa.strength = h->plane(I_STRENGTH);
            ice_strength_kernel<<<grid2d(h), 128, 0, s>>>(sa);
            ++L.n;
        }
        {   // evp_prep2 + ice_strength masking (:260-301)
            Prep2Args a;
            a.g = h->g; a.dt = dt; a.a_min = cfg.a_min; a.m_min = cfg.m_min; a.cosw = cfg.cosw; a.sinw = cfg.sinw;
            a.gravit = cfg.gravit; a.tilt_mode = cfg.tilt_mode; a.strength_masked = cfg.strength_masked;
            // Rothrock strength from the device: list semantics (0 outside icetmask == 1, ice_mechred.F90:2175)
            if (h->strength_from_itd && h->itd_cfg.kstrength == 1) a.strength_masked = 1;
            a.aiu = h->plane(W_AIU); a.umass = h->plane(W_UMASS); a.fcor = h->plane(G_FCOR);
            a.uocn = h->plane(I_UOCN); a.vocn = h->plane(I_VOCN); a.strairx = h->plane(O_STRAIRX);
            a.strairy = h->plane(O_STRAIRY); a.ss_tltx = h->plane(I_SSTLTX); a.ss_tlty = h->plane(I_SSTLTY);
            a.icetmask = h->plane(W_ICETMASK); a.strength_in = h->plane(I_STRENGTH);
            a.rhow = cfg.rhow; a.Cw = h->plane(I_CW); a.vrelc = h->plane(W_VRELC);
            a.revised = cfg.revised_evp;
            a.umask = h->iplane(1); a.iceumask = h->iplane(2);
            a.fm = h->plane(S_FM); a.strtltx = h->plane(O_STRTLTX); a.strtlty = h->plane(O_STRTLTY);
            a.strocnx = h->plane(O_STROCNX); a.strocny = h->plane(O_STROCNY);
            a.strintx = h->plane(O_STRINTX); a.strinty = h->plane(O_STRINTY);
            a.umassdti = h->plane(W_UMASSDTI); a.waterx = h->plane(W_WATERX); a.watery = h->plane(W_WATERY);
            a.forcex = h->plane(W_FORCEX); a.forcey = h->plane(W_FORCEY); a.strength = h->plane(W_STRENGTH);
            for (int q = 0; q < 12; ++q) {
                a.sA[q] = h->plane((cur ? S_STRB : S_STRA) + q);
                a.sB[q] = h->plane((cur ? S_STRA : S_STRB) + q);
            }
            a.uA = h->plane(cur ? S_UB : S_UA); a.vA = h->plane(cur ? S_VB : S_VA);
            a.uvel_init = h->plane(O_UVELINIT); a.vvel_init = h->plane(O_VVELINIT);
            a.mask = h->mask(); a.ldm = h->ldm; a.counts = h->counts();
            prep2_kernel<<<grid2d(h), 128, 0, s>>>(a);
            ++L.n;
        }
        if (cfg.kdyn == 2) {   // structure tensor back to isotropic where there is no ice (ice_dyn_eap.F90:274-289)
            EapArgs e;
            fill_eap_args(h, e);
            eap_reset_kernel<<<dim3((h->nxb + 31) / 32, h->nyb), 128, 0, s>>>(e, h->nxb, h->nyb, h->ld, h->ldm,
                                                                             h->plane(W_ICETMASK), h->mask());
            ++L.n;
        }
        if (h->use_tma) {   // per-strip row activity for the TMA producer
            dim3 grid(h->tstrips, (h->g.ny + 4 + 3) / 4);
            tile_flags_kernel<<<grid, 128, 0, s>>>(h->mask(), h->ldm, h->g.nxl, h->g.ny, h->tstrips, h->tflags(),
                                                  h->tflags_ld, h->gflags());
            ItemListLoop lp;
            memset(&lp, 0, sizeof(lp));
            if (loop_mode(h)) {
                const int nr = h->cfg.nranks;
                lp.natural = 1;
                lp.misc = h->lflags();
                lp.tflag = h->lflags() + 16 + 2 * (size_t)h->t_chunks;
                // my edge strips are the east ghost tiles of my west neighbour and the west ghost tiles of my east one
                if (nr > 1 && !h->west_closed()) lp.pub_w = h->lflags(h->west()) + 16 + h->t_chunks;
                if (nr > 1 && !h->east_closed()) lp.pub_e = h->lflags(h->east()) + 16;
            }
            item_list_kernel<<<1, kSortThreads, 0, s>>>(h->tflags(), h->tflags_ld, h->tstrips, h->t_chunks, h->t_rpc,
                                                h->g.ny,
                                                h->g.tripole ? 1 : 0, h->west_closed() ? 0 : 1, h->east_closed() ? 0 : 1,
                                                h->item_list(), h->item_count(), h->queue(0), 3 * h->cfg.ndte, lp);
            L.n += 2;
        }
        CK(cudaGetLastError());
        return EVP_OK;
    });
    B.halo(W_STRENGTH, 0, +1.0, 0.0);                          // :311
    B.halo(Ucur, 1, -1.0, 0.0, kFirstPair3D);                  // :314 (fld2: the 3-D routine)
    B.halo(Vcur, 1, -1.0, 0.0, kFirstPair3D);
    B.op([h, pb, Ucur, Uoth](Launcher &L) {
        // the other ping-pong copy must carry the same ghost ring where the loop never writes
        // (south row, closed E-W edges, open north row): start it as a copy of the current one
        CK(cudaMemcpyAsync(h->plane(Uoth), h->plane(Ucur), pb * 2, cudaMemcpyDeviceToDevice, L.s()));
        return EVP_OK;
    });
    B.sync();      // neighbours' in-loop ghost stores into that copy must come after this memcpy
    B.op([h](Launcher &L) { CK(cudaEventRecord(h->ev[1], L.s())); return EVP_OK; });

    // ---- subcycling (:336-410) ----
    P.loop_begin = P.ops.size();
    int par = cur;
    if (loop_mode(h)) {
        // subcycles 1 .. ndte-1 in the persistent kernel; the last one -- which also writes divu, shear, rdg_*,
        // prs_sig, strint* (ice_dyn_evp.F90:665-677) -- as a launch of the LAST instantiation of the
        // per-subcycle kernel (evp_loop_tma.cuh says why)
        if (cfg.ndte > 1) {
            add_loop(B, par, cfg.ndte - 1);
            par ^= (cfg.ndte - 1) & 1;
        }
        add_subcycle(B, par, true, cfg.ndte);
        par ^= 1;
    } else {
        for (int ksub = 1; ksub <= cfg.ndte; ++ksub) {
            add_subcycle(B, par, ksub == cfg.ndte, ksub);
            par ^= 1;
        }
    }
    P.loop_end = P.ops.size();
    P.end_parity = par;
    B.op([h](Launcher &L) { CK(cudaEventRecord(h->ev[2], L.s())); return EVP_OK; });

    // ---- tripole stress symmetrisation (:416-481) ----
    if (cfg.grid_tripole && cfg.kdyn != 2) {      // eap has no ice_HaloUpdate_stress
        const int S = par ? S_STRB : S_STRA;
        for (int f = 0; f < 3; ++f) {
            const int b = S + 4 * f;
            B.halo_stress(b + 0, b + 2); B.halo_stress(b + 2, b + 0);
            B.halo_stress(b + 1, b + 3); B.halo_stress(b + 3, b + 1);
        }
    }
    B.op([h, par](Launcher &L) {   // evp_finish (:490)
        FinishArgs a;
        a.g = h->g; a.rhow = h->cfg.rhow; a.cosw = h->cfg.cosw; a.sinw = h->cfg.sinw;
        a.u = h->plane(par ? S_UB : S_UA); a.v = h->plane(par ? S_VB : S_VA);
        a.uocn = h->plane(I_UOCN); a.vocn = h->plane(I_VOCN); a.aiu = h->plane(W_AIU); a.fm = h->plane(S_FM);
        a.Cw = h->plane(I_CW); a.iceumask = h->iplane(2);
        a.strocnx = h->plane(O_STROCNX); a.strocny = h->plane(O_STROCNY);
        a.work1 = h->plane(W_1); a.work2 = h->plane(W_2);
        finish_kernel<<<grid2d(h), 128, 0, L.s()>>>(a);
        ++L.n;
        return EVP_OK;
    });
    B.halo(W_1, 1, -1.0, 0.0);                                 // u2tgrid_vector (:505-506)
    B.halo(W_2, 1, -1.0, 0.0);
    B.op([h](Launcher &L) {
        cudaStream_t s = L.s();
        to_tgrid2_kernel<<<grid2d(h), 128, 0, s>>>(h->g, h->plane(G_TAREA), h->plane(G_UAREA), h->plane(W_1),
                                                  h->plane(W_2), h->plane(O_STROCNXT), h->plane(O_STROCNYT));
        ++L.n;
        double_to_int_kernel<<<(unsigned)((h->plane_elems + 255) / 256), 256, 0, s>>>(h->plane(W_ICETMASK),
                                                                                      h->iplane(3), h->plane_elems);
        ++L.n;
        CK(cudaGetLastError());
        CK(cudaEventRecord(h->ev[3], s));
        CK(cudaMemcpyAsync(h->h_counts, h->counts(), 16, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(&h->h_error, h->dev_error(), sizeof(int), cudaMemcpyDeviceToHost, s));
        return EVP_OK;
    });
}

int exec_range(Launcher &L, const Program &P, size_t b, size_t e)
{
    for (size_t i = b; i < e; ++i) {
        const Op &op = P.ops[i];
        if (op.sync) { TRY(launch_signal(L)); TRY(launch_wait(L)); }
        else TRY(op.fn(L));
    }
    return EVP_OK;
}

// the ndte loop: straight launches, or one CUDA graph per starting parity, captured on first use
int exec_loop(Launcher &L, const Program &P, long long &n_loop)
{
    evp_handle *h = L.h;
    const long long n0 = L.n;
    if (!h->cfg.use_cuda_graph) {
        TRY(exec_range(L, P, P.loop_begin, P.loop_end));
        n_loop = L.n - n0;
        return EVP_OK;
    }
    cudaGraphExec_t &ge = h->loop_graph[h->parity];
    if (!ge) {
        cudaGraph_t graph = nullptr;
        CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
        int rc = exec_range(L, P, P.loop_begin, P.loop_end);
        cudaError_t e = cudaStreamEndCapture(h->stream, &graph);
        if (rc != EVP_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
        CK(e);
        CK(cudaGraphInstantiate(&ge, graph, 0));
        CK(cudaGraphDestroy(graph));
        h->stats.loop_kernel_launches = L.n - n0;
    } else {
        L.n += h->stats.loop_kernel_launches;
    }
    CK(cudaGraphLaunch(ge, h->stream));
    n_loop = h->stats.loop_kernel_launches;
    return EVP_OK;
}

int finish_compute(evp_handle *h, const Program &P, long long n_total, long long n_loop)
{
    CK(cudaStreamSynchronize(h->stream));
    h->parity = P.end_parity;
    float ms;
    CK(cudaEventElapsedTime(&ms, h->ev[0], h->ev[1])); h->stats.prep_ms = ms;
    CK(cudaEventElapsedTime(&ms, h->ev[1], h->ev[2])); h->stats.loop_ms = ms;
    CK(cudaEventElapsedTime(&ms, h->ev[2], h->ev[3])); h->stats.finish_ms = ms;
    CK(cudaEventElapsedTime(&ms, h->ev[0], h->ev[3])); h->stats.total_ms = ms;
    h->stats.active_t_cells = (long long)h->h_counts[0];
    h->stats.active_u_cells = (long long)h->h_counts[1];
    h->stats.kernel_launches = n_total;
    h->stats.loop_kernel_launches = n_loop;
    if (h->h_error) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "halo wait timed out: a peer rank did not arrive");
        return EVP_ERR_PEER;
    }
    return EVP_OK;
}

}  // namespace

extern "C" {

int evp_gpu_compute(evp_handle_t *h, double dt)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->grid_set || !h->uploaded || !h->peers_set) return EVP_ERR_STATE;
    if (h->cfg.kdyn == 2 && !h->eap_tables_set) return EVP_ERR_STATE;
    if (!(dt > 0.0)) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    Program P;
    build_program(h, dt, P);
    Launcher L{h};
    long long n_loop = 0;
    TRY(exec_range(L, P, 0, P.loop_begin));
    TRY(exec_loop(L, P, n_loop));
    TRY(exec_range(L, P, P.loop_end, P.ops.size()));
    CK(cudaGetLastError());
    return finish_compute(h, P, L.n, n_loop);
}

// Several ranks driven by ONE host thread on ONE stream (all handles on the same device, or
// tests with fewer GPUs than ranks): the steps of all ranks are interleaved phase by phase, so
// every signal is enqueued before any wait of the same barrier and no kernel ever spins.
int evp_gpu_compute_group(evp_handle_t **hs, int n, double dt)
{
    if (!hs || n < 1 || n > kMaxRanks) return EVP_ERR_ARG;
    for (int r = 0; r < n; ++r) {
        if (!hs[r] || hs[r]->cfg.nranks != n || hs[r]->cfg.rank != r) return EVP_ERR_ARG;
        if (!hs[r]->grid_set || !hs[r]->uploaded || !hs[r]->peers_set) return EVP_ERR_STATE;
        if (hs[r]->cfg.kdyn == 2 && !hs[r]->eap_tables_set) return EVP_ERR_STATE;
        if (hs[r]->cfg.device != hs[0]->cfg.device) return EVP_ERR_UNSUPPORTED;
    }
    if (!(dt > 0.0)) return EVP_ERR_ARG;
    CK(cudaSetDevice(hs[0]->cfg.device));
    std::vector<Program> P(n);
    std::vector<Launcher> L;
    for (int r = 0; r < n; ++r) {
        CK(cudaStreamSynchronize(hs[r]->stream));
        hs[r]->stream = hs[0]->own_stream;             // one shared stream
        hs[r]->shared_stream = n > 1;
        build_program(hs[r], dt, P[r]);
        L.push_back(Launcher{hs[r]});
    }
    int rc = EVP_OK;
    for (int r = 0; r < n; ++r)
        if (P[r].ops.size() != P[0].ops.size()) rc = EVP_ERR_STATE;
    long long n_before = 0, n_after = 0;
    for (size_t i = 0; i < P[0].ops.size() && rc == EVP_OK; ++i) {
        if (i == P[0].loop_begin) n_before = L[0].n;
        if (i == P[0].loop_end) n_after = L[0].n;
        if (P[0].ops[i].sync) {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_signal(L[r]);
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_wait(L[r]);
        } else {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = P[r].ops[i].fn(L[r]);
        }
    }
    for (int r = 0; r < n; ++r) {
        int rc2 = (rc == EVP_OK) ? finish_compute(hs[r], P[r], L[r].n, n_after - n_before) : rc;
        hs[r]->stream = hs[r]->own_stream;
        hs[r]->shared_stream = false;
        if (rc2 != EVP_OK) rc = rc2;
    }
    return rc;
}

int evp_gpu_run(evp_handle_t *h, double dt, const evp_in_t *in, evp_inout_t *io, evp_out_t *out)
{
    if (!h || !in || !io) return EVP_ERR_ARG;
    // resident protocol: the state goes up only when the device does not hold it (first call, invalidated)
    const bool state_up = !h->resident || !h->state_valid || !h->uploaded;
    TRY(evp_gpu_upload(h, in, state_up ? io : nullptr));
    TRY(evp_gpu_compute(h, dt));
    return download_groups(h, io, out, h->resident ? h->dl_mask : (unsigned)EVP_DL_ALL);
}

}  // extern "C"

namespace {
// the steps of restartfile's stress fix-up (source/ice_restart_driver.F90:370-394, 431-466) for one handle
void build_restart_fixup(evp_handle *h, std::vector<Op> &ops)
{
    Builder B{h, ops};
    const int S = h->parity ? S_STRB : S_STRA;
    B.sync();                                                    // every rank's state is uploaded
    if (h->cfg.grid_tripole) {                                   // :370-394
        for (int f = 0; f < 3; ++f) {
            const int b = S + 4 * f;
            B.halo_stress(b + 0, b + 2); B.halo_stress(b + 2, b + 0);
            B.halo_stress(b + 1, b + 3); B.halo_stress(b + 3, b + 1);
        }
    }
    B.op([h, S](Launcher &L) {                                   // :431-466
        dim3 grid((h->nxb + h->nyb + 127) / 128, 12);
        zero_sw_ghosts_kernel<<<grid, 128, 0, L.s()>>>(h->plane(S), h->plane_elems, h->nxb, h->nyb, h->ld);
        ++L.n;
        CK(cudaGetLastError());
        return EVP_OK;
    });
    B.sync();                                                    // inbox buffers are free again
}
int finish_fixup(evp_handle *h)
{
    CK(cudaMemcpyAsync(&h->h_error, h->dev_error(), sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (h->h_error) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "halo wait timed out: a peer rank did not arrive");
        return EVP_ERR_PEER;
    }
    return EVP_OK;
}
}  // namespace

extern "C" {

int evp_gpu_restart_fixup(evp_handle_t *h)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->grid_set || !h->uploaded || !h->peers_set) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    std::vector<Op> ops;
    build_restart_fixup(h, ops);
    Launcher L{h};
    for (const Op &op : ops) {
        if (op.sync) { TRY(launch_signal(L)); TRY(launch_wait(L)); }
        else TRY(op.fn(L));
    }
    return finish_fixup(h);
}

// all ranks of a one-device group on one stream, steps interleaved (see evp_gpu_compute_group)
int evp_gpu_restart_fixup_group(evp_handle_t **hs, int n)
{
    if (!hs || n < 1 || n > kMaxRanks) return EVP_ERR_ARG;
    for (int r = 0; r < n; ++r) {
        if (!hs[r] || hs[r]->cfg.nranks != n || hs[r]->cfg.rank != r) return EVP_ERR_ARG;
        if (!hs[r]->grid_set || !hs[r]->uploaded || !hs[r]->peers_set) return EVP_ERR_STATE;
        if (hs[r]->cfg.device != hs[0]->cfg.device) return EVP_ERR_UNSUPPORTED;
    }
    CK(cudaSetDevice(hs[0]->cfg.device));
    std::vector<std::vector<Op>> P(n);
    std::vector<Launcher> L;
    for (int r = 0; r < n; ++r) CK(cudaStreamSynchronize(hs[r]->stream));      // (before any stream is switched: an early
    for (int r = 0; r < n; ++r) {                                                // return leaves every handle as it was)
        hs[r]->stream = hs[0]->own_stream;
        build_restart_fixup(hs[r], P[r]);
        L.push_back(Launcher{hs[r]});
    }
    int rc = EVP_OK;
    for (size_t i = 0; i < P[0].size() && rc == EVP_OK; ++i) {
        if (P[0][i].sync) {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_signal(L[r]);
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_wait(L[r]);
        } else {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = P[r][i].fn(L[r]);
        }
    }
    for (int r = 0; r < n; ++r) {
        const int rc2 = rc == EVP_OK ? finish_fixup(hs[r]) : rc;
        hs[r]->stream = hs[r]->own_stream;
        if (rc2 != EVP_OK) rc = rc2;
    }
    return rc;
}

}  // extern "C"

// ---- horizontal transport, first stage: departure points from the resident velocities -------------------
namespace {
// device time of a transport stage: events around its kernels and around its ghost-cell updates
void stage_mark(Builder &B, int k)
{
    evp_handle *h = B.h;
    B.op([h, k](Launcher &L) {
        if (!h->tev[k]) CK(cudaEventCreate(&h->tev[k]));
        CK(cudaEventRecord(h->tev[k], L.s()));
        return EVP_OK;
    });
}
int stage_times(evp_handle *h, int slot, int ev)
{
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, h->tev[ev], h->tev[ev + 1])); h->tr_ms[slot] = ms;
    CK(cudaEventElapsedTime(&ms, h->tev[ev + 1], h->tev[ev + 2])); h->tr_ms[slot + 1] = ms;
    return EVP_OK;
}

// departure_points (source/ice_transport_remap.F90:1493-1670) into the work planes W_1 (dpx), W_2 (dpy), then the two
// ghost-cell updates of horizontal_remap (:574-578)
void build_departure(evp_handle *h, double dt, int midpt, std::vector<Op> &ops)
{
    Builder B{h, ops};
    B.sync();                                                    // every rank has left its previous operation
    stage_mark(B, 5);
    B.op([h, dt, midpt](Launcher &L) {
        DepartArgs a;
        a.nxb = h->nxb; a.nyb = h->nyb; a.ld = h->ld; a.dt = dt; a.midpt = midpt;
        a.u = h->plane(h->parity ? S_UB : S_UA); a.v = h->plane(h->parity ? S_VB : S_VA);
        a.dxu = h->tr; a.dyu = h->tr + h->plane_elems; a.htn = h->tr + 2 * h->plane_elems; a.hte = h->tr + 3 * h->plane_elems;
        a.dpx = h->plane(W_1); a.dpy = h->plane(W_2); a.stop = h->tr_stop;
        CK(cudaMemsetAsync(h->tr_stop, 0, sizeof(int), L.s()));
        departure_first_kernel<<<grid2d(h), 128, 0, L.s()>>>(a);
        departure_midpt_kernel<<<grid2d(h), 128, 0, L.s()>>>(a);
        L.n += 2;
        CK(cudaGetLastError());
        return EVP_OK;
    });
    stage_mark(B, 6);
    B.halo(W_1, 1, -1.0, 0.0);                                   // ice_HaloUpdate(dpx, NEcorner, vector)
    B.halo(W_2, 1, -1.0, 0.0);
    B.sync();                                                    // inbox buffers are free again
    stage_mark(B, 7);
}
int finish_departure(evp_handle *h)
{
    CK(cudaMemcpyAsync(&h->h_tr_stop, h->tr_stop, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&h->h_error, h->dev_error(), sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (h->h_error) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "halo wait timed out: a peer rank did not arrive");
        return EVP_ERR_PEER;
    }
    return stage_times(h, 0, 5);
}
}  // namespace

extern "C" {

int evp_gpu_set_transport_grid(evp_handle_t *h, const evp_transport_grid_t *g)
{
    if (!h || !g || !g->dxu || !g->dyu || !g->HTN || !g->HTE) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    if (!h->tr) {
        CK(cudaMalloc(&h->tr, 4 * h->plane_elems * sizeof(double) + 256));
        CK(cudaMemset(h->tr, 0, 4 * h->plane_elems * sizeof(double) + 256));
        h->tr_stop = reinterpret_cast<int *>(h->tr + 4 * h->plane_elems);
    }
    const double *src[4] = {g->dxu, g->dyu, g->HTN, g->HTE};
    for (int q = 0; q < 4; ++q) TRY(h2d_any(h, h->tr + (size_t)q * h->plane_elems, src[q], 0));
    TRY(h2d_flush(h));
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

int evp_gpu_departure_points_compute(evp_handle_t *h, double dt, int l_dp_midpt)
{
    if (!h || !(dt > 0.0)) return EVP_ERR_ARG;
    if (!h->tr || !h->uploaded || !h->peers_set) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    std::vector<Op> ops;
    build_departure(h, dt, l_dp_midpt ? 1 : 0, ops);
    Launcher L{h};
    for (const Op &op : ops) {
        if (op.sync) { TRY(launch_signal(L)); TRY(launch_wait(L)); }
        else TRY(op.fn(L));
    }
    return finish_departure(h);
}

// all ranks of a one-device group on one stream, steps interleaved (see evp_gpu_compute_group)
int evp_gpu_departure_points_group(evp_handle_t **hs, int n, double dt, int l_dp_midpt)
{
    if (!hs || n < 1 || n > kMaxRanks || !(dt > 0.0)) return EVP_ERR_ARG;
    for (int r = 0; r < n; ++r) {
        if (!hs[r] || hs[r]->cfg.nranks != n || hs[r]->cfg.rank != r) return EVP_ERR_ARG;
        if (!hs[r]->tr || !hs[r]->uploaded || !hs[r]->peers_set) return EVP_ERR_STATE;
        if (hs[r]->cfg.device != hs[0]->cfg.device) return EVP_ERR_UNSUPPORTED;
    }
    CK(cudaSetDevice(hs[0]->cfg.device));
    std::vector<std::vector<Op>> P(n);
    std::vector<Launcher> L;
    for (int r = 0; r < n; ++r) CK(cudaStreamSynchronize(hs[r]->stream));      // (before any stream is switched: an early
    for (int r = 0; r < n; ++r) {                                                // return leaves every handle as it was)
        hs[r]->stream = hs[0]->own_stream;
        build_departure(hs[r], dt, l_dp_midpt ? 1 : 0, P[r]);
        L.push_back(Launcher{hs[r]});
    }
    int rc = EVP_OK;
    for (size_t i = 0; i < P[0].size() && rc == EVP_OK; ++i) {
        if (P[0][i].sync) {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_signal(L[r]);
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_wait(L[r]);
        } else {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = P[r][i].fn(L[r]);
        }
    }
    for (int r = 0; r < n; ++r) {
        const int rc2 = rc == EVP_OK ? finish_departure(hs[r]) : rc;
        hs[r]->stream = hs[r]->own_stream;
        if (rc2 != EVP_OK) rc = rc2;
    }
    return rc;
}

int evp_gpu_departure_points_fetch(evp_handle_t *h, double *dpx, double *dpy, int *l_stop, int *istop, int *jstop)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->tr) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    D2HBatch B{h, {}};
    h->batch_dst.clear();
    TRY(B.r8(W_1, dpx));
    TRY(B.r8(W_2, dpy));
    TRY(B.flush());
    CK(cudaStreamSynchronize(h->stream));
    const int code = h->h_tr_stop;                               // 1 + (j-1)*nxb + (i-1), Fortran i, j of the block
    if (l_stop) *l_stop = code != 0;
    if (istop) *istop = code ? (code - 1) % h->nxb + 1 : 0;
    if (jstop) *jstop = code ? (code - 1) / h->nxb + 1 : 0;
    return EVP_OK;
}

int evp_gpu_departure_points(evp_handle_t *h, double dt, int l_dp_midpt, double *dpx, double *dpy, int *l_stop, int *istop,
                             int *jstop)
{
    TRY(evp_gpu_departure_points_compute(h, dt, l_dp_midpt));
    return evp_gpu_departure_points_fetch(h, dpx, dpy, l_stop, istop, jstop);
}

// ---- horizontal transport, field reconstruction: make_masks + construct_fields + the ghost-cell updates ----------
}  // extern "C"
namespace {
// plane k of the reconstruction arena
struct RcLayout {
    int ncat, ntrace;
    size_t geom(int k) const { return (size_t)k; }                                   // hm xav yav xxav yyav
    size_t mm(int n) const { return 5 + (size_t)n; }
    size_t tm(int k) const { return 5 + (size_t)(ncat + 1) + k; }                       // k = (n-1)*ntrace + nt
    size_t mass(int f, int n) const { return tm(ncat * ntrace) + (size_t)f * (ncat + 1) + n; }           // f: mc mx my
    size_t trc(int f, int k) const { return mass(3, 0) + (size_t)f * ncat * ntrace + k; }                // f: tc tx ty
    size_t mflx(int e) const { return trc(3, 0) + (size_t)e * (ncat + 1 + (size_t)ncat * ntrace); }      // e: east north
    size_t mtflx(int e) const { return mflx(e) + (ncat + 1); }
    size_t planes() const { return mflx(2); }
};
RcLayout rc_layout(const evp_handle *h) { return RcLayout{h->rc_args.ncat, h->rc_args.ntrace}; }

// reconstruct_kernel, then ice_HaloUpdate of mc (center, scalar), mx, my (center, vector) and tc, tx, ty
// (source/ice_transport_remap.F90:579-584, :616-621; maskhalo_remap = .false.)
void build_reconstruct(evp_handle *h, std::vector<Op> &ops)
{
    Builder B{h, ops};
    const RcLayout l = rc_layout(h);
    B.sync();
    stage_mark(B, 0);
    B.op([h, l](Launcher &L) {
        int nparents = 0;
        for (int nt = 0; nt < l.ntrace; ++nt) if (h->rc_args.slot[nt] >= nparents) nparents = h->rc_args.slot[nt] + 1;
        const dim3 tiles((h->nxb + kRcDirectW - 1) / kRcDirectW, (h->nyb + kRcDirectH - 1) / kRcDirectH, l.ncat + 1);
        reconstruct_kernel<<<tiles, dim3(kRcDirectW, kRcDirectH), (size_t)nparents * 128 * 18, L.s()>>>(h->rc_args, nparents);
        ++L.n;
        CK(cudaGetLastError());
        return EVP_OK;
    });
    stage_mark(B, 1);
    auto pl = [h](size_t k) { return h->rc + k * h->plane_elems; };
    {   // mc | mx my | tc | tx ty are consecutive planes: one stack, the gradients change sign across the fold
        const int nm = l.ncat + 1, nt = l.ncat * l.ntrace;
        const NegRanges neg{{nm, 3 * nm + nt}, {3 * nm, 3 * nm + 3 * nt}};
        B.halo_multi(pl(l.mass(0, 0)), 3 * nm + 3 * nt, h->plane_elems, neg, 0.0);
    }
    B.sync();
    stage_mark(B, 2);
}
int finish_ops(evp_handle *h)
{
    CK(cudaMemcpyAsync(&h->h_error, h->dev_error(), sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (h->h_error) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "halo wait timed out: a peer rank did not arrive");
        return EVP_ERR_PEER;
    }
    return stage_times(h, 2, 0);
}
}  // namespace
extern "C" {

int evp_gpu_get_transport_timing(evp_handle_t *h, double ms[6])
{
    if (!h || !ms) return EVP_ERR_ARG;
    for (int k = 0; k < 6; ++k) ms[k] = h->tr_ms[k];
    return EVP_OK;
}

int evp_gpu_set_transport_tracers(evp_handle_t *h, int ncat, int ntrace, const int *tracer_type, const int *depend,
                                  const int *has_dependents, const double *hm, const double *xav, const double *yav,
                                  const double *xxav, const double *yyav)
{
    if (!h || ncat < 1 || ncat > 255 || ntrace < 0 || !hm || !xav || !yav || !xxav || !yyav) return EVP_ERR_ARG;
    if (ntrace > 0 && (!tracer_type || !depend || !has_dependents)) return EVP_ERR_ARG;
    if (ntrace > kRemapMaxTrace) return EVP_ERR_TOO_LARGE;
    ReconArgs a;
    memset(&a, 0, sizeof(a));
    int parents = 0;
    for (int nt = 0; nt < ntrace; ++nt) {
        const int ty = tracer_type[nt], d = depend[nt];
        if (ty < 1 || ty > 3) return EVP_ERR_ARG;
        a.ttype[nt] = (signed char)ty;
        a.slot[nt] = -1;
        if (ty != 1) {
            // the reference reads the sums of tracer depend(nt) when it reaches tracer nt: the parent of a type-2 tracer
            // is an earlier type-1 tracer, that of a type-3 tracer an earlier type-2 tracer, both flagged has_dependents
            // (init_transport, ice_transport_driver.F90:90-118)
            if (d < 1 || d > nt || tracer_type[d - 1] != ty - 1 || !has_dependents[d - 1]) return EVP_ERR_ARG;
            a.parent[nt] = (unsigned char)(d - 1);
        }
        if (has_dependents[nt]) {
            if (ty == 3) return EVP_ERR_ARG;
            if (parents == kRemapMaxParents) return EVP_ERR_TOO_LARGE;
            a.slot[nt] = (signed char)parents++;
        }
    }
    {   // tracer numbers type by type
        int k = 0;
        for (int ty = 1; ty <= 3; ++ty) {
            for (int nt = 0; nt < ntrace; ++nt) if (tracer_type[nt] == ty) a.order[k++] = (unsigned char)nt;
            if (ty == 1) a.n1 = (short)k;
            if (ty == 2) a.n2 = (short)(k - a.n1);
        }
    }
    CK(cudaSetDevice(h->cfg.device));
    const RcLayout l{ncat, ntrace};
    if (h->rc && (h->rc_args.ncat != ncat || h->rc_args.ntrace != ntrace)) {
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaFree(h->rc));
        h->rc = nullptr;
        if (h->st) { CK(cudaFree(h->st)); h->st = nullptr; }
        h->st_loaded = h->st_done = h->rc_remapped = false;
    }
    if (!h->rc) {
        CK(cudaMalloc(&h->rc, l.planes() * h->plane_elems * sizeof(double) + 1024));
        CK(cudaMemset(h->rc, 0, l.planes() * h->plane_elems * sizeof(double) + 1024));
        h->rc_stop = reinterpret_cast<int *>(h->rc + l.planes() * h->plane_elems);        // [ncat + 1] <= 256 words
    }
    auto pl = [h](size_t k) { return h->rc + k * h->plane_elems; };
    a.nxb = h->nxb; a.nyb = h->nyb; a.ld = h->ld; a.pe = h->plane_elems; a.ncat = ncat; a.ntrace = ntrace;
    a.hm = pl(0); a.xav = pl(1); a.yav = pl(2); a.xxav = pl(3); a.yyav = pl(4);
    a.mm = pl(l.mm(0)); a.tm = pl(l.tm(0));
    a.mc = pl(l.mass(0, 0)); a.mx = pl(l.mass(1, 0)); a.my = pl(l.mass(2, 0));
    a.tc = pl(l.trc(0, 0)); a.tx = pl(l.trc(1, 0)); a.ty = pl(l.trc(2, 0));
    h->rc_args = a;
    h->rc_loaded = false;
    const double *src[5] = {hm, xav, yav, xxav, yyav};
    for (int k = 0; k < 5; ++k) TRY(h2d_any(h, pl(k), src[k], 0));
    TRY(h2d_flush(h));
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

int evp_gpu_construct_fields_upload(evp_handle_t *h, const double *mm, const double *tm)
{
    if (!h || !mm) return EVP_ERR_ARG;
    if (!h->rc) return EVP_ERR_STATE;
    const RcLayout l = rc_layout(h);
    if (l.ntrace > 0 && !tm) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    const size_t cells = (size_t)h->nxb * h->nyb;
    auto pl = [h](size_t k) { return h->rc + k * h->plane_elems; };
    for (int n = 0; n <= l.ncat; ++n) TRY(h2d_any(h, pl(l.mm(n)), mm + (size_t)n * cells, 0));
    for (int k = 0; k < l.ncat * l.ntrace; ++k) TRY(h2d_any(h, pl(l.tm(k)), tm + (size_t)k * cells, 0));
    TRY(h2d_flush(h));
    CK(cudaStreamSynchronize(h->stream));                          // the host arrays may be reused
    h->rc_loaded = true;
    return EVP_OK;
}

int evp_gpu_construct_fields_compute(evp_handle_t *h)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->rc || !h->rc_loaded || !h->peers_set) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    std::vector<Op> ops;
    build_reconstruct(h, ops);
    Launcher L{h};
    for (const Op &op : ops) {
        if (op.sync) { TRY(launch_signal(L)); TRY(launch_wait(L)); }
        else TRY(op.fn(L));
    }
    return finish_ops(h);
}

// all ranks of a one-device group on one stream, steps interleaved (see evp_gpu_compute_group)
int evp_gpu_construct_fields_group(evp_handle_t **hs, int n)
{
    if (!hs || n < 1 || n > kMaxRanks) return EVP_ERR_ARG;
    for (int r = 0; r < n; ++r) {
        if (!hs[r] || hs[r]->cfg.nranks != n || hs[r]->cfg.rank != r) return EVP_ERR_ARG;
        if (!hs[r]->rc || !hs[r]->rc_loaded || !hs[r]->peers_set) return EVP_ERR_STATE;
        if (hs[r]->cfg.device != hs[0]->cfg.device) return EVP_ERR_UNSUPPORTED;
        if (hs[r]->rc_args.ncat != hs[0]->rc_args.ncat || hs[r]->rc_args.ntrace != hs[0]->rc_args.ntrace) return EVP_ERR_ARG;
    }
    CK(cudaSetDevice(hs[0]->cfg.device));
    std::vector<std::vector<Op>> P(n);
    std::vector<Launcher> L;
    for (int r = 0; r < n; ++r) CK(cudaStreamSynchronize(hs[r]->stream));      // (before any stream is switched: an early
    for (int r = 0; r < n; ++r) {                                                // return leaves every handle as it was)
        hs[r]->stream = hs[0]->own_stream;
        build_reconstruct(hs[r], P[r]);
        L.push_back(Launcher{hs[r]});
    }
    int rc = EVP_OK;
    for (size_t i = 0; i < P[0].size() && rc == EVP_OK; ++i) {
        if (P[0][i].sync) {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_signal(L[r]);
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_wait(L[r]);
        } else {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = P[r][i].fn(L[r]);
        }
    }
    for (int r = 0; r < n; ++r) {
        const int rc2 = rc == EVP_OK ? finish_ops(hs[r]) : rc;
        hs[r]->stream = hs[r]->own_stream;
        if (rc2 != EVP_OK) rc = rc2;
    }
    return rc;
}

int evp_gpu_construct_fields_fetch(evp_handle_t *h, double *mc, double *mx, double *my, double *tc, double *tx, double *ty)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->rc || !h->rc_loaded) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    const RcLayout l = rc_layout(h);
    const size_t cells = (size_t)h->nxb * h->nyb;
    auto pl = [h](size_t k) { return h->rc + k * h->plane_elems; };
    D2HBatch B{h, {}};
    h->batch_dst.clear();
    double *mass[3] = {mc, mx, my}, *trc[3] = {tc, tx, ty};
    for (int f = 0; f < 3; ++f) {
        if (mass[f]) for (int n = 0; n <= l.ncat; ++n) TRY(B.add(pl(l.mass(f, n)), mass[f] + (size_t)n * cells, 0));
        if (trc[f]) for (int k = 0; k < l.ncat * l.ntrace; ++k) TRY(B.add(pl(l.trc(f, k)), trc[f] + (size_t)k * cells, 0));
    }
    TRY(B.flush());
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

int evp_gpu_construct_fields(evp_handle_t *h, const double *mm, const double *tm, double *mc, double *mx, double *my,
                             double *tc, double *tx, double *ty)
{
    TRY(evp_gpu_construct_fields_upload(h, mm, tm));
    TRY(evp_gpu_construct_fields_compute(h));
    return evp_gpu_construct_fields_fetch(h, mc, mx, my, tc, tx, ty);
}

// ---- all of horizontal_remap on the device ----------------------------------------------------------------------
}  // extern "C"
namespace {
int rc_slots(const evp_handle *h)
{
    int n = 0;
    for (int nt = 0; nt < h->rc_args.ntrace; ++nt) if (h->rc_args.slot[nt] >= n) n = h->rc_args.slot[nt] + 1;
    return n;
}
// horizontal_remap (source/ice_transport_remap.F90:309-850): reconstruction and departure points with their ghost-cell
// updates (the two programs above), then the fluxes through the east and north edges and update_fields of every category
void build_remap(evp_handle *h, double dt, std::vector<Op> &ops)
{
    { Builder B0{h, ops}; stage_mark(B0, 3); }
    build_reconstruct(h, ops);
    build_departure(h, dt, 1, ops);
    Builder B{h, ops};
    const RcLayout l = rc_layout(h);
    B.op([h, l](Launcher &L) {
        auto pl = [h](size_t k) { return h->rc + k * h->plane_elems; };
        const int nslots = rc_slots(h);
        FluxArgs f;
        f.a = h->rc_args;
        f.dpx = h->plane(W_1); f.dpy = h->plane(W_2);
        f.dxu = h->tr; f.dyu = h->tr + h->plane_elems;
        for (int e = 0; e < 2; ++e) { f.mflx[e] = pl(l.mflx(e)); f.mtflx[e] = pl(l.mtflx(e)); }
        // shared memory per block: (1 + ntrace + 3 slots) doubles x 128 threads (above 48 KB: opt in; at most 177 KB)
        const size_t fsm = (size_t)(1 + l.ntrace + 3 * nslots) * 128 * sizeof(double);
        if (fsm > 48 * 1024) CK(cudaFuncSetAttribute(edge_flux_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsm));
        edge_flux_kernel<<<dim3(2 * ((h->nxb + 127) / 128), h->nyb, l.ncat + 1), 128, fsm, L.s()>>>(f, nslots);
        CK(cudaGetLastError());
        UpdateArgs u;
        u.a = h->rc_args;
        u.tarear = h->plane(G_TAREAR);
        for (int e = 0; e < 2; ++e) { u.mflx[e] = pl(l.mflx(e)); u.mtflx[e] = pl(l.mtflx(e)); }
        u.mm = pl(l.mm(0)); u.tm = pl(l.tm(0)); u.stop = h->rc_stop;
        CK(cudaMemsetAsync(h->rc_stop, 0, (l.ncat + 1) * sizeof(int), L.s()));
        update_fields_kernel<<<dim3((h->nxb + 127) / 128, h->nyb, l.ncat + 1), 128, (size_t)nslots * 128 * 16, L.s()>>>(u, nslots);
        CK(cudaGetLastError());
        L.n += 2;
        return EVP_OK;
    });
    stage_mark(B, 4);
}
int finish_remap(evp_handle *h)
{
    const int ncat = h->rc_args.ncat;
    CK(cudaMemcpyAsync(&h->h_tr_stop, h->tr_stop, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(h->h_rc_stop, h->rc_stop, (ncat + 1) * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&h->h_error, h->dev_error(), sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (h->h_error) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "halo wait timed out: a peer rank did not arrive");
        return EVP_ERR_PEER;
    }
    h->rc_loaded = false;                                        // mm, tm now hold the transported state
    h->rc_remapped = true;
    float ms = 0.f;
    TRY(stage_times(h, 2, 0));
    TRY(stage_times(h, 0, 5));
    CK(cudaEventElapsedTime(&ms, h->tev[7], h->tev[4])); h->tr_ms[4] = ms;
    CK(cudaEventElapsedTime(&ms, h->tev[3], h->tev[4])); h->tr_ms[5] = ms;
    return EVP_OK;
}
int remap_ready(const evp_handle *h)
{
    if (!h->tr || !h->rc || !h->rc_loaded || !h->uploaded || !h->peers_set || !h->grid_set) return EVP_ERR_STATE;
    return EVP_OK;
}
}  // namespace
extern "C" {

int evp_gpu_horizontal_remap_compute(evp_handle_t *h, double dt)
{
    if (!h || !(dt > 0.0)) return EVP_ERR_ARG;
    TRY(remap_ready(h));
    CK(cudaSetDevice(h->cfg.device));
    std::vector<Op> ops;
    build_remap(h, dt, ops);
    Launcher L{h};
    for (const Op &op : ops) {
        if (op.sync) { TRY(launch_signal(L)); TRY(launch_wait(L)); }
        else TRY(op.fn(L));
    }
    return finish_remap(h);
}

// all ranks of a one-device group on one stream, steps interleaved (see evp_gpu_compute_group)
int evp_gpu_horizontal_remap_group(evp_handle_t **hs, int n, double dt)
{
    if (!hs || n < 1 || n > kMaxRanks || !(dt > 0.0)) return EVP_ERR_ARG;
    for (int r = 0; r < n; ++r) {
        if (!hs[r] || hs[r]->cfg.nranks != n || hs[r]->cfg.rank != r) return EVP_ERR_ARG;
        TRY(remap_ready(hs[r]));
        if (hs[r]->cfg.device != hs[0]->cfg.device) return EVP_ERR_UNSUPPORTED;
        if (hs[r]->rc_args.ncat != hs[0]->rc_args.ncat || hs[r]->rc_args.ntrace != hs[0]->rc_args.ntrace) return EVP_ERR_ARG;
    }
    CK(cudaSetDevice(hs[0]->cfg.device));
    std::vector<std::vector<Op>> P(n);
    std::vector<Launcher> L;
    for (int r = 0; r < n; ++r) CK(cudaStreamSynchronize(hs[r]->stream));      // (before any stream is switched: an early
    for (int r = 0; r < n; ++r) {                                                // return leaves every handle as it was)
        hs[r]->stream = hs[0]->own_stream;
        build_remap(hs[r], dt, P[r]);
        L.push_back(Launcher{hs[r]});
    }
    int rc = EVP_OK;
    for (size_t i = 0; i < P[0].size() && rc == EVP_OK; ++i) {
        if (P[0][i].sync) {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_signal(L[r]);
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_wait(L[r]);
        } else {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = P[r][i].fn(L[r]);
        }
    }
    for (int r = 0; r < n; ++r) {
        const int rc2 = rc == EVP_OK ? finish_remap(hs[r]) : rc;
        hs[r]->stream = hs[r]->own_stream;
        if (rc2 != EVP_OK) rc = rc2;
    }
    return rc;
}

int evp_gpu_horizontal_remap_fetch(evp_handle_t *h, double *mm, double *tm, int *stop, int *stop_cat, int *istop, int *jstop)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->rc || !h->rc_remapped) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    const RcLayout l = rc_layout(h);
    const size_t cells = (size_t)h->nxb * h->nyb;
    auto pl = [h](size_t k) { return h->rc + k * h->plane_elems; };
    D2HBatch B{h, {}};
    h->batch_dst.clear();
    if (mm) for (int n = 0; n <= l.ncat; ++n) TRY(B.add(pl(l.mm(n)), mm + (size_t)n * cells, 0));
    if (tm) for (int k = 0; k < l.ncat * l.ntrace; ++k) TRY(B.add(pl(l.tm(k)), tm + (size_t)k * cells, 0));
    TRY(B.flush());
    CK(cudaStreamSynchronize(h->stream));
    // what the reference would report before abort_ice: bad departure points (:552-563), then the first category (open
    // water first) whose new area is negative (:783-825)
    int kind = 0, cat = 0, code = 0;
    if (h->h_tr_stop) { kind = 1; code = h->h_tr_stop; }
    else
        for (int n = 0; n <= l.ncat; ++n)
            if (h->h_rc_stop[n]) { kind = n == 0 ? 2 : 3; cat = n; code = h->h_rc_stop[n]; break; }
    if (stop) *stop = kind;
    if (stop_cat) *stop_cat = cat;
    if (istop) *istop = code ? (code - 1) % h->nxb + 1 : 0;
    if (jstop) *jstop = code ? (code - 1) / h->nxb + 1 : 0;
    return EVP_OK;
}

int evp_gpu_horizontal_remap(evp_handle_t *h, double dt, double *mm, double *tm, int *stop, int *stop_cat, int *istop, int *jstop)
{
    TRY(evp_gpu_construct_fields_upload(h, mm, tm));
    TRY(evp_gpu_horizontal_remap_compute(h, dt));
    return evp_gpu_horizontal_remap_fetch(h, mm, tm, stop, stop_cat, istop, jstop);
}

// ---- transport_remap: the caller of horizontal_remap, from and to the ice state -------------------------------------
}  // extern "C"
namespace {
struct StLayout {
    int ncat, ntrcr;
    size_t aice0() const { return 0; }
    size_t aicen(int n) const { return 1 + (size_t)n; }                                // n = 0 .. ncat-1
    size_t vicen(int n) const { return 1 + (size_t)ncat + n; }
    size_t vsnon(int n) const { return 1 + 2 * (size_t)ncat + n; }
    size_t trcrn(int k) const { return 1 + 3 * (size_t)ncat + k; }                      // k = (n-1)*ntrcr + it-1
    size_t planes() const { return trcrn(ncat * ntrcr); }
};
StLayout st_layout(const evp_handle *h) { return StLayout{h->rc_args.ncat, h->rc_args.ntrace - 2}; }

// transport_remap (source/ice_transport_driver.F90:198-627; l_conservation_check = l_monotonicity_check = .false.):
// state_to_tracers, horizontal_remap, tracers_to_state, bound_state (source/ice_state.F90:173-238)
void build_transport(evp_handle *h, double dt, int nt_qsno, int nslyr, double rhos, double Lfresh, std::vector<Op> &ops)
{
    const RcLayout l = rc_layout(h);
    const StLayout sl = st_layout(h);
    auto args = [h, l, sl, nt_qsno, nslyr, rhos, Lfresh]() {
        StateArgs s;
        auto sp = [h](size_t k) { return h->st + k * h->plane_elems; };
        s.nxb = h->nxb; s.nyb = h->nyb; s.ld = h->ld; s.pe = h->plane_elems;
        s.ncat = sl.ncat; s.ntrcr = sl.ntrcr; s.nt_qsno = nt_qsno; s.nslyr = nslyr; s.rhosLf = rhos * Lfresh;
        s.aice0 = sp(sl.aice0()); s.aicen = sp(sl.aicen(0)); s.vicen = sp(sl.vicen(0)); s.vsnon = sp(sl.vsnon(0));
        s.trcrn = sp(sl.trcrn(0));
        s.aim = h->rc + l.mm(0) * h->plane_elems; s.trm = h->rc + l.tm(0) * h->plane_elems;
        return s;
    };
    {
        Builder B{h, ops};
        B.op([h, l, args](Launcher &L) {
            state_to_tracers_kernel<<<dim3((h->nxb + 127) / 128, h->nyb, l.ncat + 1), 128, 0, L.s()>>>(args());
            ++L.n;
            CK(cudaGetLastError());
            return EVP_OK;
        });
    }
    build_remap(h, dt, ops);
    Builder B{h, ops};
    B.op([h, l, args](Launcher &L) {
        tracers_to_state_kernel<<<dim3((h->nxb + 127) / 128, h->nyb, l.ncat + 1), 128, 0, L.s()>>>(args());
        ++L.n;
        CK(cudaGetLastError());
        return EVP_OK;
    });
    auto sp = [h](size_t k) { return h->st + k * h->plane_elems; };
    B.sync();
    // bound_state: aicen | vicen | vsnon | trcrn are consecutive planes, all center / scalar
    B.halo_multi(sp(sl.aicen(0)), 3 * sl.ncat + sl.ncat * sl.ntrcr, h->plane_elems, NegRanges{{0, 0}, {0, 0}}, 0.0);
    B.sync();
    stage_mark(B, 4);
}
int transport_ready(const evp_handle *h)
{
    if (!h->tr || !h->rc || !h->st || !h->st_loaded || !h->uploaded || !h->peers_set || !h->grid_set) return EVP_ERR_STATE;
    if (h->rc_args.ntrace < 2) return EVP_ERR_STATE;             // hice, hsno lead the tracer table
    return EVP_OK;
}
int finish_transport(evp_handle *h)
{
    h->rc_loaded = true;                                         // state_to_tracers filled mm, tm
    TRY(finish_remap(h));
    h->st_loaded = false;
    h->st_done = true;
    return EVP_OK;
}
}  // namespace
extern "C" {

int evp_gpu_transport_remap_upload(evp_handle_t *h, const double *aice0, const double *aicen, const double *trcrn,
                                   const double *vicen, const double *vsnon)
{
    if (!h || !aice0 || !aicen || !vicen || !vsnon) return EVP_ERR_ARG;
    if (!h->rc || h->rc_args.ntrace < 2) return EVP_ERR_STATE;
    const StLayout sl = st_layout(h);
    if (sl.ntrcr > 0 && !trcrn) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    if (!h->st) {
        CK(cudaMalloc(&h->st, sl.planes() * h->plane_elems * sizeof(double)));
        CK(cudaMemset(h->st, 0, sl.planes() * h->plane_elems * sizeof(double)));
    }
    const size_t cells = (size_t)h->nxb * h->nyb;
    auto sp = [h](size_t k) { return h->st + k * h->plane_elems; };
    TRY(h2d_any(h, sp(sl.aice0()), aice0, 0));
    for (int n = 0; n < sl.ncat; ++n) {
        TRY(h2d_any(h, sp(sl.aicen(n)), aicen + (size_t)n * cells, 0));
        TRY(h2d_any(h, sp(sl.vicen(n)), vicen + (size_t)n * cells, 0));
        TRY(h2d_any(h, sp(sl.vsnon(n)), vsnon + (size_t)n * cells, 0));
    }
    for (int k = 0; k < sl.ncat * sl.ntrcr; ++k) TRY(h2d_any(h, sp(sl.trcrn(k)), trcrn + (size_t)k * cells, 0));
    TRY(h2d_flush(h));
    CK(cudaStreamSynchronize(h->stream));
    h->st_loaded = true;
    return EVP_OK;
}

int evp_gpu_transport_remap_compute(evp_handle_t *h, double dt, int nt_qsno, int nslyr, double rhos, double Lfresh)
{
    if (!h || !(dt > 0.0) || nslyr < 0) return EVP_ERR_ARG;
    TRY(transport_ready(h));
    CK(cudaSetDevice(h->cfg.device));
    std::vector<Op> ops;
    build_transport(h, dt, nt_qsno, nslyr, rhos, Lfresh, ops);
    Launcher L{h};
    for (const Op &op : ops) {
        if (op.sync) { TRY(launch_signal(L)); TRY(launch_wait(L)); }
        else TRY(op.fn(L));
    }
    return finish_transport(h);
}

// all ranks of a one-device group on one stream, steps interleaved (see evp_gpu_compute_group)
int evp_gpu_transport_remap_group(evp_handle_t **hs, int n, double dt, int nt_qsno, int nslyr, double rhos, double Lfresh)
{
    if (!hs || n < 1 || n > kMaxRanks || !(dt > 0.0) || nslyr < 0) return EVP_ERR_ARG;
    for (int r = 0; r < n; ++r) {
        if (!hs[r] || hs[r]->cfg.nranks != n || hs[r]->cfg.rank != r) return EVP_ERR_ARG;
        TRY(transport_ready(hs[r]));
        if (hs[r]->cfg.device != hs[0]->cfg.device) return EVP_ERR_UNSUPPORTED;
        if (hs[r]->rc_args.ncat != hs[0]->rc_args.ncat || hs[r]->rc_args.ntrace != hs[0]->rc_args.ntrace) return EVP_ERR_ARG;
    }
    CK(cudaSetDevice(hs[0]->cfg.device));
    std::vector<std::vector<Op>> P(n);
    std::vector<Launcher> L;
    for (int r = 0; r < n; ++r) CK(cudaStreamSynchronize(hs[r]->stream));      // (before any stream is switched: an early
    for (int r = 0; r < n; ++r) {                                                // return leaves every handle as it was)
        hs[r]->stream = hs[0]->own_stream;
        build_transport(hs[r], dt, nt_qsno, nslyr, rhos, Lfresh, P[r]);
        L.push_back(Launcher{hs[r]});
    }
    int rc = EVP_OK;
    for (size_t i = 0; i < P[0].size() && rc == EVP_OK; ++i) {
        if (P[0][i].sync) {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_signal(L[r]);
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = launch_wait(L[r]);
        } else {
            for (int r = 0; r < n && rc == EVP_OK; ++r) rc = P[r][i].fn(L[r]);
        }
    }
    for (int r = 0; r < n; ++r) {
        const int rc2 = rc == EVP_OK ? finish_transport(hs[r]) : rc;
        hs[r]->stream = hs[r]->own_stream;
        if (rc2 != EVP_OK) rc = rc2;
    }
    return rc;
}

int evp_gpu_transport_remap_fetch(evp_handle_t *h, double *aice0, double *aicen, double *trcrn, double *vicen, double *vsnon,
                                  int *stop, int *stop_cat, int *istop, int *jstop)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->st || !h->st_done) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    const StLayout sl = st_layout(h);
    const size_t cells = (size_t)h->nxb * h->nyb;
    auto sp = [h](size_t k) { return h->st + k * h->plane_elems; };
    D2HBatch B{h, {}};
    h->batch_dst.clear();
    if (aice0) TRY(B.add(sp(sl.aice0()), aice0, 0));
    for (int n = 0; n < sl.ncat; ++n) {
        if (aicen) TRY(B.add(sp(sl.aicen(n)), aicen + (size_t)n * cells, 0));
        if (vicen) TRY(B.add(sp(sl.vicen(n)), vicen + (size_t)n * cells, 0));
        if (vsnon) TRY(B.add(sp(sl.vsnon(n)), vsnon + (size_t)n * cells, 0));
    }
    if (trcrn) for (int k = 0; k < sl.ncat * sl.ntrcr; ++k) TRY(B.add(sp(sl.trcrn(k)), trcrn + (size_t)k * cells, 0));
    TRY(B.flush());
    CK(cudaStreamSynchronize(h->stream));
    return evp_gpu_horizontal_remap_fetch(h, nullptr, nullptr, stop, stop_cat, istop, jstop);
}

int evp_gpu_transport_remap(evp_handle_t *h, double dt, int nt_qsno, int nslyr, double rhos, double Lfresh, double *aice0,
                            double *aicen, double *trcrn, double *vicen, double *vsnon, int *stop, int *stop_cat, int *istop,
                            int *jstop)
{
    TRY(evp_gpu_transport_remap_upload(h, aice0, aicen, trcrn, vicen, vsnon));
    TRY(evp_gpu_transport_remap_compute(h, dt, nt_qsno, nslyr, rhos, Lfresh));
    return evp_gpu_transport_remap_fetch(h, aice0, aicen, trcrn, vicen, vsnon, stop, stop_cat, istop, jstop);
}

int evp_gpu_principal_stress(evp_handle_t *h, double *sig1, double *sig2)
{
    if (!h || !sig1 || !sig2) return EVP_ERR_ARG;
    if (!h->uploaded) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    const int S = h->parity ? S_STRB : S_STRA;          // the copy that holds the current stresses
    principal_stress_kernel<<<grid2d(h), 128, 0, h->stream>>>(h->nxb, h->nyb, h->ld, h->plane(S), h->plane(S + 4),
                                                            h->plane(S + 8), h->plane(O_PRSSIG), h->plane(W_1),
                                                            h->plane(W_2));
    CK(cudaGetLastError());
    D2HBatch B{h, {}};
    h->batch_dst.clear();
    TRY(B.r8(W_1, sig1));
    TRY(B.r8(W_2, sig2));
    TRY(B.flush());
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

// ---- EAP (kdyn = 2): lookup tables, structure tensor, diagnostics ---------------------------------------------------
int evp_gpu_set_eap_tables(evp_handle_t *h, const evp_eap_tables_t *t)
{
    if (!h || !t || !t->s11r || !t->s12r || !t->s22r || !t->s11s || !t->s12s || !t->s22s) return EVP_ERR_ARG;
    if (h->cfg.kdyn != 2) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    if (!h->eap) {
        const size_t elems = 6 * (size_t)kEapTablePitch + (size_t)kEapPlanes * h->plane_elems;
        CK(cudaMalloc(&h->eap, elems * sizeof(double)));
        CK(cudaMemsetAsync(h->eap, 0, elems * sizeof(double), h->stream));
        // init_eap's state (source/ice_dyn_eap.F90:531-548): a11_k = p5, a12_k = c0, the diagnostics zero
        fill_plane_kernel<<<(unsigned)((4 * h->plane_elems + 255) / 256), 256, 0, h->stream>>>(h->eap_plane(0), 0.5,
                                                                                            4 * h->plane_elems);
        CK(cudaGetLastError());
    }
    const double *src[6] = {t->s11r, t->s12r, t->s22r, t->s11s, t->s12s, t->s22s};
    for (int q = 0; q < 6; ++q)
        CK(cudaMemcpyAsync(h->eap + (size_t)q * kEapTablePitch, src[q], kEapTableElems * sizeof(double),
                           cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->eap_tables_set = true;
    return EVP_OK;
}

int evp_gpu_upload_eap(evp_handle_t *h, const evp_eap_state_t *st)
{
    if (!h || !st) return EVP_ERR_ARG;
    if (h->cfg.kdyn != 2 || !h->eap) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    for (int q = 0; q < 4; ++q) {
        if (st->a11[q]) TRY(h2d_any(h, h->eap_plane(q), st->a11[q], 0));
        if (st->a12[q]) TRY(h2d_any(h, h->eap_plane(4 + q), st->a12[q], 0));
    }
    if (st->a11_avg) TRY(h2d_any(h, h->eap_plane(8), st->a11_avg, 0));
    if (st->a12_avg) TRY(h2d_any(h, h->eap_plane(9), st->a12_avg, 0));
    TRY(h2d_flush(h));
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

int evp_gpu_download_eap(evp_handle_t *h, evp_eap_state_t *st, evp_eap_out_t *out)
{
    if (!h) return EVP_ERR_ARG;
    if (h->cfg.kdyn != 2 || !h->eap) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    D2HBatch B{h, {}};
    h->batch_dst.clear();
    if (st) {
        for (int q = 0; q < 4; ++q) {
            TRY(B.add(h->eap_plane(q), st->a11[q], 0));
            TRY(B.add(h->eap_plane(4 + q), st->a12[q], 0));
        }
        TRY(B.add(h->eap_plane(8), st->a11_avg, 0));
        TRY(B.add(h->eap_plane(9), st->a12_avg, 0));
    }
    if (out) {
        double *dst[9] = {out->e11, out->e12, out->e22, out->s11, out->s12, out->s22, out->yieldstress11,
                          out->yieldstress12, out->yieldstress22};
        for (int q = 0; q < 9; ++q) TRY(B.add(h->eap_plane(10 + q), dst[q], 0));
    }
    TRY(B.flush());
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

}  // extern "C"

// ---- evp_gpu_principal_stress_static: principal_stress without any allocation ---------------------------
// Scratch of the existing entry point: two work planes of the handle's arena + the staging area (cudaMalloc at init /
// first use).  Here: one file-scope device array -- 4 input planes + 2 output planes of at most
// EVP_STATIC_MAX_CELLS cells, dense -- that every GPU which loads the library holds until exit.
// All calls share it (it keeps the last call's contents; every element a call reads is written by that call first,
// nothing has to start at zero); calls that overlap on one GPU race.
static __device__ __align__(256) double g_static_scratch[6 * (size_t)EVP_STATIC_MAX_CELLS];

namespace {
constexpr int kStaticMaxDevices = 64;
double *g_static_ptr[kStaticMaxDevices] = {nullptr};       // device address of g_static_scratch, per device

// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: resolve it once per
// device, with that device current.
int static_scratch_ptr(int device, double **out)
{
    if (device < 0 || device >= kStaticMaxDevices) return EVP_ERR_NO_DEVICE;
    if (!g_static_ptr[device]) {
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_static_scratch));
        g_static_ptr[device] = static_cast<double *>(p);
    }
    *out = g_static_ptr[device];
    return EVP_OK;
}
}  // namespace

extern "C" {

int evp_gpu_principal_stress_static(int device, int nx_block, int ny_block, const double *stressp_1,
                                    const double *stressm_1, const double *stress12_1, const double *prs_sig,
                                    double *sig1, double *sig2)
{
    if (!stressp_1 || !stressm_1 || !stress12_1 || !prs_sig || !sig1 || !sig2) return EVP_ERR_ARG;
    if (nx_block < 1 || ny_block < 1) return EVP_ERR_ARG;
    const long long cells = (long long)nx_block * ny_block;
    if (cells > (long long)EVP_STATIC_MAX_CELLS) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "evp_gpu_principal_stress_static: %d x %d = %lld cells exceed the static "
                 "scratch of %d cells", nx_block, ny_block, cells, (int)EVP_STATIC_MAX_CELLS);
        return EVP_ERR_TOO_LARGE;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1 || device < 0 || device >= ndev) return EVP_ERR_NO_DEVICE;
    CK(cudaSetDevice(device));
    double *scr = nullptr;
    TRY(static_scratch_ptr(device, &scr));
    const size_t n = (size_t)cells, bytes = n * sizeof(double);
    const double *src[4] = {stressp_1, stressm_1, stress12_1, prs_sig};
    for (int q = 0; q < 4; ++q) CK(cudaMemcpy(scr + q * n, src[q], bytes, cudaMemcpyHostToDevice));
    dim3 grid((nx_block + 127) / 128, ny_block);
    // legacy default stream: no stream is created; the blocking cudaMemcpy calls around the launch order everything
    principal_stress_kernel<<<grid, 128, 0, 0>>>(nx_block, ny_block, nx_block, scr, scr + n, scr + 2 * n, scr + 3 * n,
                                                 scr + 4 * n, scr + 5 * n);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(0));
    CK(cudaMemcpy(sig1, scr + 4 * n, bytes, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(sig2, scr + 5 * n, bytes, cudaMemcpyDeviceToHost));
    return EVP_OK;
}

int evp_gpu_get_trace(evp_handle_t *h, unsigned long long *buf, int n)
{
    if (!h || !buf) return EVP_ERR_ARG;
    if (!h->trace_on) return EVP_ERR_STATE;
    const int have = h->cfg.ndte * 8;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaMemcpy(buf, h->arena + h->off_trace, sizeof(unsigned long long) * (size_t)(n < have ? n : have),
                  cudaMemcpyDeviceToHost));
    return EVP_OK;
}

int evp_gpu_peer_blob_size(void) { return (int)sizeof(PeerBlob); }

int evp_gpu_get_peer_blob(evp_handle_t *h, void *blob)
{
    if (!h || !blob) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    PeerBlob b;
    memset(&b, 0, sizeof(b));
    CK(cudaIpcGetMemHandle(&b.arena, h->arena));
    b.rank = h->cfg.rank; b.device = h->cfg.device; b.pid = (long long)getpid();
    b.arena_ptr = (unsigned long long)(uintptr_t)h->arena;
    memcpy(blob, &b, sizeof(b));
    return EVP_OK;
}

int evp_gpu_set_peers(evp_handle_t *h, const void *blobs)
{
    if (!h || !blobs) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    const PeerBlob *b = static_cast<const PeerBlob *>(blobs);
    for (int r = 0; r < h->cfg.nranks; ++r) {
        if (r == h->cfg.rank) continue;
        if (b[r].rank != r) return EVP_ERR_PEER;
        void *p = nullptr;
        if (b[r].pid == (long long)getpid()) {          // same process: plain pointer (+ peer access)
            if (b[r].device != h->cfg.device) {
                cudaError_t e2 = cudaDeviceEnablePeerAccess(b[r].device, 0);
                if (e2 != cudaSuccess && e2 != cudaErrorPeerAccessAlreadyEnabled) {
                    snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaDeviceEnablePeerAccess(%d): %s", b[r].device,
                             cudaGetErrorString(e2));
                    return EVP_ERR_PEER;
                }
                cudaGetLastError();
            }
            h->peer_arena[r] = reinterpret_cast<char *>((uintptr_t)b[r].arena_ptr);
            continue;
        }
        cudaError_t e = cudaIpcOpenMemHandle(&p, b[r].arena, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
            return EVP_ERR_PEER;
        }
        h->peer_arena[r] = static_cast<char *>(p);
        h->peer_opened[r] = true;
    }
    h->peers_set = true;
    return EVP_OK;
}

}  // extern "C"
