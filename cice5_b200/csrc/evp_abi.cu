// evp_abi.cu -- host side of the C ABI in include/evp_b200.h: the B200-native replacement of
// CICE5's `evp(dt)` driver (source/ice_dyn_evp.F90:68-510).  All arithmetic runs in the CUDA
// kernels of evp_kernels.cuh / evp_prep.cuh; there is no CPU fallback.
#include "../../include/evp_b200.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "evp_kernels.cuh"
#include "evp_prep.cuh"

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

// ---- plane directory ----------------------------------------------------------------------
enum Plane {
    // grid (evp_grid_t order)
    G_DXT, G_DYT, G_DXHY, G_DYHX, G_CXP, G_CYP, G_CXM, G_CYM, G_TAREA, G_UAREA, G_TAREAR, G_UAREAR,
    G_TINYAREA, G_FCOR,
    // per-call inputs (evp_in_t order)
    I_AICE, I_VICE, I_VSNO, I_AICE_INIT, I_STRAIRXT, I_STRAIRYT, I_UOCN, I_VOCN, I_SSTLTX, I_SSTLTY,
    I_CW, I_STRENGTH,
    // dynamic state, ping-pong A / B
    S_UA, S_VA, S_UB, S_VB,
    S_STRA, S_STRA_END = S_STRA + 11,
    S_STRB, S_STRB_END = S_STRB + 11,
    S_FM,
    // evp locals (ice_dyn_evp.F90:120-137)
    W_TMASS, W_UMASS, W_AIU, W_UMASSDTI, W_WATERX, W_WATERY, W_FORCEX, W_FORCEY, W_STRENGTH,
    W_ICETMASK, W_1, W_2,
    // outputs (evp_out_t order, first 17)
    O_STRAIRX, O_STRAIRY, O_STRTLTX, O_STRTLTY, O_STROCNX, O_STROCNY, O_STRINTX, O_STRINTY,
    O_STROCNXT, O_STROCNYT, O_DIVU, O_SHEAR, O_RDGCONV, O_RDGSHEAR, O_PRSSIG, O_UVELINIT, O_VVELINIT,
    NPLANES
};

constexpr int kInboxRows = 4;

struct PeerBlob {                 // exchanged between the tasks of one node
    cudaIpcMemHandle_t arena;
    int rank, device, pid_tag;
};

}  // namespace

struct evp_handle {
    evp_config_t cfg;
    Geom g;
    Consts k;
    int nxb, nyb, ld, ldm;
    size_t plane_elems;           // doubles per plane (ld * nyb)
    // one arena, identical layout on every rank, so a peer's plane is peer_base + offset
    char *arena = nullptr;
    size_t arena_bytes = 0;
    size_t off_planes = 0, off_inbox = 0, off_flags = 0, off_int = 0, off_mask = 0, off_counts = 0;
    char *peer_arena[kMaxRanks] = {nullptr};
    bool peer_opened[kMaxRanks] = {false};
    bool grid_set = false, uploaded = false, peers_set = false;
    int parity = 0;               // which ping-pong copy holds the current state (0 = A)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaGraphExec_t loop_graph = nullptr;
    evp_stats_t stats;
    unsigned long long h_counts[2];

    double *plane(int p, int rank = -1) const
    {
        char *base = (rank < 0 || rank == cfg.rank) ? arena : peer_arena[rank];
        return reinterpret_cast<double *>(base + off_planes) + (size_t)p * plane_elems;
    }
    double *inbox(int rank = -1) const
    {
        char *base = (rank < 0 || rank == cfg.rank) ? arena : peer_arena[rank];
        return reinterpret_cast<double *>(base + off_inbox);
    }
    int *iplane(int p) const   // 0 tmask, 1 umask, 2 iceumask, 3 icetmask(int)
    {
        return reinterpret_cast<int *>(arena + off_int) + (size_t)p * plane_elems;
    }
    unsigned char *mask() const { return reinterpret_cast<unsigned char *>(arena + off_mask); }
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

// set_evp_parameters, source/ice_dyn_shared.F90:209-257 (classic branch); constants
// drivers/cice/ice_constants.F90:142-148 are computed, not literals.
void set_evp_parameters(const evp_config_t &c, Consts &k)
{
    const double c1 = 1.0, c2 = 2.0, p5 = 0.5, eyc = 0.36;
    k.p166 = c1 / 6.0; k.p333 = c1 / 3.0; k.p111 = c1 / 9.0;
    k.p055 = k.p111 * p5; k.p027 = k.p055 * p5; k.p222 = c2 / 9.0;
    const double dte = c.dt_init / (double)c.ndte;
    const double dtei = c1 / dte;
    k.ecci = 0.25;
    const double tdamp2 = c2 * eyc * c.dt_init;
    const double dte2T = dte / tdamp2;
    k.revp = 0.0;
    k.arlx1i = dte2T;
    k.brlx = c.dt_init * dtei;
    k.denom1 = c1 / (c1 + k.arlx1i);
    k.cosw = c.cosw; k.sinw = c.sinw; k.rhow = c.rhow;
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
            return "evp_gpu: unsupported configuration (needs kdyn=1, revised_evp=.false., "
                   "tripole only with cyclic E-W, nranks <= 8 dividing nx_global)";
        case EVP_ERR_CUDA: return "evp_gpu: CUDA runtime error (see evp_gpu_last_cuda_error)";
        case EVP_ERR_STATE: return "evp_gpu: call order violated (init -> set_grid -> [set_peers] -> run)";
        case EVP_ERR_NO_DEVICE: return "evp_gpu: no usable CUDA device; this library has no CPU fallback";
        case EVP_ERR_PEER: return "evp_gpu: multi-GPU peer mapping failed";
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
    if (cfg->kdyn != 1 || cfg->revised_evp != 0) return EVP_ERR_UNSUPPORTED;
    if (cfg->nranks < 1 || cfg->nranks > kMaxRanks || cfg->rank < 0 || cfg->rank >= cfg->nranks)
        return EVP_ERR_UNSUPPORTED;
    if (cfg->nx_global < 2 || cfg->ny_global < 3 || cfg->ndte < 1 || !(cfg->dt_init > 0.0)) return EVP_ERR_ARG;
    if (cfg->nx_global % cfg->nranks) return EVP_ERR_UNSUPPORTED;
    if (cfg->nx_block != cfg->nx_global / cfg->nranks + 2 || cfg->ny_block != cfg->ny_global + 2)
        return EVP_ERR_ARG;
    if (cfg->nx_global / cfg->nranks < 2) return EVP_ERR_UNSUPPORTED;
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
    set_evp_parameters(*cfg, h->k);
    h->nxb = cfg->nx_block; h->nyb = cfg->ny_block;
    h->ld = h->nxb; h->ldm = h->nxb;
    h->plane_elems = (size_t)h->ld * h->nyb;
    Geom &g = h->g;
    g.nxl = h->nxb - 2; g.ny = h->nyb - 2; g.ld = h->ld; g.nx_global = cfg->nx_global;
    g.lo = cfg->rank * g.nxl; g.nranks = cfg->nranks; g.rank = cfg->rank;
    g.tripole = cfg->ns_boundary == EVP_NS_TRIPOLE; g.ew_cyclic = cfg->ew_boundary == EVP_EW_CYCLIC;
    g.inbox_ld = (int)align_up((size_t)h->nxb, 16);

    size_t off = 0;
    h->off_planes = off; off = align_up(off + (size_t)NPLANES * h->plane_elems * sizeof(double), 256);
    h->off_inbox = off;  off = align_up(off + (size_t)kInboxRows * g.inbox_ld * sizeof(double), 256);
    h->off_flags = off;  off = align_up(off + 4096, 256);
    h->off_int = off;    off = align_up(off + 4 * h->plane_elems * sizeof(int), 256);
    h->off_mask = off;   off = align_up(off + (size_t)h->ldm * h->nyb, 256);
    h->off_counts = off; off = align_up(off + 64, 256);
    h->arena_bytes = off;
    cudaError_t e = cudaMalloc(&h->arena, h->arena_bytes);
    if (e != cudaSuccess) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaMalloc(%zu bytes): %s", h->arena_bytes, cudaGetErrorString(e));
        delete h;
        return EVP_ERR_CUDA;
    }
    CK(cudaMemset(h->arena, 0, h->arena_bytes));
    CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    for (auto &ev : h->ev) CK(cudaEventCreate(&ev));
    memset(&h->stats, 0, sizeof(h->stats));
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
    if (h->loop_graph) cudaGraphExecDestroy(h->loop_graph);
    for (int r = 0; r < kMaxRanks; ++r)
        if (h->peer_opened[r]) cudaIpcCloseMemHandle(h->peer_arena[r]);
    for (auto &ev : h->ev) if (ev) cudaEventDestroy(ev);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->arena) cudaFree(h->arena);
    delete h;
    return EVP_OK;
}

int evp_gpu_synchronize(evp_handle_t *h)
{
    if (!h) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    return EVP_OK;
}

static int h2d(evp_handle *h, int p, const double *src)
{
    if (!src) return EVP_ERR_ARG;
    CK(cudaMemcpyAsync(h->plane(p), src, h->plane_elems * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    return EVP_OK;
}
static int d2h(evp_handle *h, int p, double *dst)
{
    if (!dst) return EVP_OK;
    CK(cudaMemcpyAsync(dst, h->plane(p), h->plane_elems * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    return EVP_OK;
}
#define TRY(x) do { int rc_ = (x); if (rc_ != EVP_OK) return rc_; } while (0)

int evp_gpu_set_grid(evp_handle_t *h, const evp_grid_t *gr)
{
    if (!h || !gr) return EVP_ERR_ARG;
    CK(cudaSetDevice(h->cfg.device));
    const double *src[14] = {gr->dxt, gr->dyt, gr->dxhy, gr->dyhx, gr->cxp, gr->cyp, gr->cxm, gr->cym,
                             gr->tarea, gr->uarea, gr->tarear, gr->uarear, gr->tinyarea, gr->fcor};
    for (int p = 0; p < 14; ++p) TRY(h2d(h, G_DXT + p, src[p]));
    if (!gr->tmask || !gr->umask) return EVP_ERR_ARG;
    CK(cudaMemcpyAsync(h->iplane(0), gr->tmask, h->plane_elems * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->iplane(1), gr->umask, h->plane_elems * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->grid_set = true;
    return EVP_OK;
}

int evp_gpu_upload(evp_handle_t *h, const evp_in_t *in, const evp_inout_t *io)
{
    if (!h || !in || !io) return EVP_ERR_ARG;
    if (!h->grid_set) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    const double *src[12] = {in->aice, in->vice, in->vsno, in->aice_init, in->strairxT, in->strairyT,
                             in->uocn, in->vocn, in->ss_tltx, in->ss_tlty, in->Cdn_ocn, in->strength};
    for (int p = 0; p < 12; ++p) TRY(h2d(h, I_AICE + p, src[p]));
    // dynamic state always lands in copy A
    TRY(h2d(h, S_UA, io->uvel));
    TRY(h2d(h, S_VA, io->vvel));
    for (int c = 0; c < 4; ++c) {
        TRY(h2d(h, S_STRA + c, io->stressp[c]));
        TRY(h2d(h, S_STRA + 4 + c, io->stressm[c]));
        TRY(h2d(h, S_STRA + 8 + c, io->stress12[c]));
    }
    if (!io->iceumask) return EVP_ERR_ARG;
    CK(cudaMemcpyAsync(h->iplane(2), io->iceumask, h->plane_elems * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    h->parity = 0;
    h->uploaded = true;
    return EVP_OK;
}

int evp_gpu_download(evp_handle_t *h, evp_inout_t *io, evp_out_t *out)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->uploaded) return EVP_ERR_STATE;
    CK(cudaSetDevice(h->cfg.device));
    if (io) {
        const int U = h->parity ? S_UB : S_UA, V = h->parity ? S_VB : S_VA;
        const int S = h->parity ? S_STRB : S_STRA;
        TRY(d2h(h, U, io->uvel));
        TRY(d2h(h, V, io->vvel));
        for (int c = 0; c < 4; ++c) {
            TRY(d2h(h, S + c, io->stressp[c]));
            TRY(d2h(h, S + 4 + c, io->stressm[c]));
            TRY(d2h(h, S + 8 + c, io->stress12[c]));
        }
        TRY(d2h(h, S_FM, io->fm));
        if (io->iceumask)
            CK(cudaMemcpyAsync(io->iceumask, h->iplane(2), h->plane_elems * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    }
    if (out) {
        double *dst[17] = {out->strairx, out->strairy, out->strtltx, out->strtlty, out->strocnx, out->strocny,
                           out->strintx, out->strinty, out->strocnxT, out->strocnyT, out->divu, out->shear,
                           out->rdg_conv, out->rdg_shear, out->prs_sig, out->uvel_init, out->vvel_init};
        for (int p = 0; p < 17; ++p) TRY(d2h(h, O_STRAIRX + p, dst[p]));
        TRY(d2h(h, W_TMASS, out->tmass));
        TRY(d2h(h, W_STRENGTH, out->strength));
        if (out->icetmask)
            CK(cudaMemcpyAsync(out->icetmask, h->iplane(3), h->plane_elems * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
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

// ice_HaloUpdate of one device plane: push my edges / top rows, then pre-fill + fold.
int do_halo(Launcher &L, int plane, int loc, double sgn, double fill)
{
    evp_handle *h = L.h;
    const Geom &g = h->g;
    PushArgs pa;
    pa.g = g;
    pa.plane = h->plane(plane);
    pa.planeW = h->west_closed() ? nullptr : h->plane(plane, h->west());
    pa.planeE = h->east_closed() ? nullptr : h->plane(plane, h->east());
    for (int r = 0; r < kMaxRanks; ++r) pa.inbox[r] = r < g.nranks ? h->inbox(r) : nullptr;
    pa.ioff = loc;              // center: source nx+1-ig, NEcorner: nx-ig
    pa.inbox_row0 = 0;
    const int nt = g.ny + (g.tripole ? g.nxl : 0);
    halo_push_kernel<<<(nt + 255) / 256, 256, 0, L.s()>>>(pa);
    ++L.n;
    const int na = g.nxl + 2 + g.ny;
    halo_apply_kernel<<<(na + 255) / 256, 256, 0, L.s()>>>(h->plane(plane), g, h->inbox(), 0, loc, sgn, fill,
                                                        h->west_closed(), h->east_closed());
    ++L.n;
    CK(cudaGetLastError());
    return EVP_OK;
}

// ice_HaloUpdate_stress(a1, a2): needs a2's top physical row mirrored with center offsets.
int do_halo_stress(Launcher &L, int a1, int a2)
{
    evp_handle *h = L.h;
    const Geom &g = h->g;
    if (g.tripole) {
        PushArgs pa;
        pa.g = g;
        pa.plane = h->plane(a2);
        pa.planeW = nullptr; pa.planeE = nullptr;     // no ordinary neighbour copies (:3346-3353)
        for (int r = 0; r < kMaxRanks; ++r) pa.inbox[r] = r < g.nranks ? h->inbox(r) : nullptr;
        pa.ioff = 0;
        pa.inbox_row0 = 0;
        const int nt = g.ny + g.nxl;
        halo_push_kernel<<<(nt + 255) / 256, 256, 0, L.s()>>>(pa);
        ++L.n;
    }
    const int na = g.nxl + 2 + g.ny + 2;
    halo_stress_apply_kernel<<<(na + 255) / 256, 256, 0, L.s()>>>(h->plane(a1), g, h->inbox(), 0,
                                                               h->west_closed(), h->east_closed());
    ++L.n;
    CK(cudaGetLastError());
    return EVP_OK;
}

void fill_loop_args(const evp_handle *h, int from_parity, LoopArgs &a)
{
    const Geom &g = h->g;
    memset(&a, 0, sizeof(a));
    a.nxl = g.nxl; a.ny = g.ny; a.ld = g.ld; a.ldm = h->ldm;
    a.rows_per_chunk = 32;
    a.nx_global = g.nx_global; a.lo = g.lo; a.nranks = g.nranks; a.rank = g.rank; a.tripole = g.tripole;
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
    if (!h->west_closed()) { a.uW = h->plane(U1, h->west()); a.vW = h->plane(V1, h->west()); }
    if (!h->east_closed()) { a.uE = h->plane(U1, h->east()); a.vE = h->plane(V1, h->east()); }
    for (int r = 0; r < g.nranks; ++r) a.inbox[r] = h->inbox(r);
    a.k = h->k;
}

// one subcycle: stress + stepu + halo pushes fused; then the tripole fold of the new u, v
int launch_subcycle(Launcher &L, int from_parity, bool last)
{
    evp_handle *h = L.h;
    const Geom &g = h->g;
    LoopArgs a;
    fill_loop_args(h, from_parity, a);
    const int strips = (g.nxl + kUColsPerWarp - 1) / kUColsPerWarp;
    dim3 grid((strips + kWarpsPerBlock - 1) / kWarpsPerBlock, (g.ny + a.rows_per_chunk - 1) / a.rows_per_chunk);
    if (last) evp_subcycle_kernel<true><<<grid, kWarpsPerBlock * 32, 0, L.s()>>>(a);
    else evp_subcycle_kernel<false><<<grid, kWarpsPerBlock * 32, 0, L.s()>>>(a);
    ++L.n;
    if (g.tripole) {
        fold_uv_kernel<<<(g.nxl + 2 + 127) / 128, 128, 0, L.s()>>>(a.u1, a.v1, g.ld, g.nxl, g.ny, g.nx_global,
                                                                 g.lo, h->inbox(), g.inbox_ld);
        ++L.n;
    }
    return EVP_OK;
}

}  // namespace

extern "C" {

int evp_gpu_compute(evp_handle_t *h, double dt)
{
    if (!h) return EVP_ERR_ARG;
    if (!h->grid_set || !h->uploaded || !h->peers_set) return EVP_ERR_STATE;
    if (!(dt > 0.0)) return EVP_ERR_ARG;
    if (h->cfg.nranks > 1) return EVP_ERR_UNSUPPORTED;   // multi-GPU sync arrives with evp_multi
    CK(cudaSetDevice(h->cfg.device));
    const Geom &g = h->g;
    const evp_config_t &cfg = h->cfg;
    Launcher L{h};
    cudaStream_t s = h->stream;
    const size_t pb = h->plane_elems * sizeof(double);

    CK(cudaEventRecord(h->ev[0], s));
    // init_history_dyn contract (ice_flux.F90:639-673) + evp's own zeroing (ice_dyn_evp.F90:174-182)
    CK(cudaMemsetAsync(h->plane(O_STRAIRX), 0, pb * 17, s));   // all output planes
    CK(cudaMemsetAsync(h->plane(S_FM), 0, pb, s));
    CK(cudaMemsetAsync(h->counts(), 0, 16, s));

    {   // evp_prep1 (:194)
        Prep1Args a;
        a.g = g; a.rhoi = cfg.rhoi; a.rhos = cfg.rhos; a.a_min = cfg.a_min; a.m_min = cfg.m_min;
        a.aice = h->plane(I_AICE); a.vice = h->plane(I_VICE); a.vsno = h->plane(I_VSNO);
        a.strairxT = h->plane(I_STRAIRXT); a.strairyT = h->plane(I_STRAIRYT); a.tmask = h->iplane(0);
        a.tmass = h->plane(W_TMASS); a.work1 = h->plane(W_1); a.work2 = h->plane(W_2);
        a.icetmask = h->plane(W_ICETMASK);
        prep1_kernel<<<grid2d(h), 128, 0, s>>>(a);
        ++L.n;
    }
    TRY(do_halo(L, W_ICETMASK, 0, +1.0, 0.0));                 // :210
    TRY(do_halo(L, W_1, 0, -1.0, 0.0));                        // t2ugrid_vector (ice_grid.F90:1817)
    TRY(do_halo(L, W_2, 0, -1.0, 0.0));
    {   // to_ugrid: umass, aiu, strairx, strairy (:218-241)
        ToUArgs a;
        a.g = g; a.tarea = h->plane(G_TAREA); a.uarea = h->plane(G_UAREA);
        a.in[0] = h->plane(W_TMASS); a.out[0] = h->plane(W_UMASS);
        a.in[1] = h->plane(I_AICE_INIT); a.out[1] = h->plane(W_AIU);
        a.in[2] = h->plane(W_1); a.out[2] = h->plane(O_STRAIRX);
        a.in[3] = h->plane(W_2); a.out[3] = h->plane(O_STRAIRY);
        to_ugrid4_kernel<<<grid2d(h), 128, 0, s>>>(a);
        ++L.n;
    }
    {   // evp_prep2 + ice_strength masking (:260-301)
        Prep2Args a;
        a.g = g; a.dt = dt; a.a_min = cfg.a_min; a.m_min = cfg.m_min; a.cosw = cfg.cosw; a.sinw = cfg.sinw;
        a.gravit = cfg.gravit; a.tilt_mode = cfg.tilt_mode; a.strength_masked = cfg.strength_masked;
        a.aiu = h->plane(W_AIU); a.umass = h->plane(W_UMASS); a.fcor = h->plane(G_FCOR);
        a.uocn = h->plane(I_UOCN); a.vocn = h->plane(I_VOCN); a.strairx = h->plane(O_STRAIRX);
        a.strairy = h->plane(O_STRAIRY); a.ss_tltx = h->plane(I_SSTLTX); a.ss_tlty = h->plane(I_SSTLTY);
        a.icetmask = h->plane(W_ICETMASK); a.strength_in = h->plane(I_STRENGTH);
        a.umask = h->iplane(1); a.iceumask = h->iplane(2);
        a.fm = h->plane(S_FM); a.strtltx = h->plane(O_STRTLTX); a.strtlty = h->plane(O_STRTLTY);
        a.strocnx = h->plane(O_STROCNX); a.strocny = h->plane(O_STROCNY);
        a.strintx = h->plane(O_STRINTX); a.strinty = h->plane(O_STRINTY);
        a.umassdti = h->plane(W_UMASSDTI); a.waterx = h->plane(W_WATERX); a.watery = h->plane(W_WATERY);
        a.forcex = h->plane(W_FORCEX); a.forcey = h->plane(W_FORCEY); a.strength = h->plane(W_STRENGTH);
        const int cur = h->parity;
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
    TRY(do_halo(L, W_STRENGTH, 0, +1.0, 0.0));                 // :311
    TRY(do_halo(L, h->parity ? S_UB : S_UA, 1, -1.0, 0.0));    // :314 (fld2)
    TRY(do_halo(L, h->parity ? S_VB : S_VA, 1, -1.0, 0.0));
    // the other copy must carry the same ghost ring where the loop never writes (south row,
    // closed E-W edges, open north row): start it as a copy of the current one
    CK(cudaMemcpyAsync(h->plane(h->parity ? S_UA : S_UB), h->plane(h->parity ? S_UB : S_UA), pb * 2,
                       cudaMemcpyDeviceToDevice, s));
    CK(cudaGetLastError());

    // ---- subcycling (:336-410) ----
    CK(cudaEventRecord(h->ev[1], s));
    const long long n_before_loop = L.n;
    for (int ksub = 1; ksub <= cfg.ndte; ++ksub) {
        TRY(launch_subcycle(L, h->parity, ksub == cfg.ndte));
        h->parity ^= 1;
    }
    CK(cudaGetLastError());
    const long long n_loop = L.n - n_before_loop;
    CK(cudaEventRecord(h->ev[2], s));

    // ---- tripole stress symmetrisation (:416-481) ----
    if (cfg.grid_tripole) {
        const int S = h->parity ? S_STRB : S_STRA;
        for (int f = 0; f < 3; ++f) {
            const int b = S + 4 * f;
            TRY(do_halo_stress(L, b + 0, b + 2)); TRY(do_halo_stress(L, b + 2, b + 0));
            TRY(do_halo_stress(L, b + 1, b + 3)); TRY(do_halo_stress(L, b + 3, b + 1));
        }
    }
    {   // evp_finish (:490) + u2tgrid_vector (:505-506)
        FinishArgs a;
        a.g = g; a.rhow = cfg.rhow; a.cosw = cfg.cosw; a.sinw = cfg.sinw;
        a.u = h->plane(h->parity ? S_UB : S_UA); a.v = h->plane(h->parity ? S_VB : S_VA);
        a.uocn = h->plane(I_UOCN); a.vocn = h->plane(I_VOCN); a.aiu = h->plane(W_AIU); a.fm = h->plane(S_FM);
        a.Cw = h->plane(I_CW); a.iceumask = h->iplane(2);
        a.strocnx = h->plane(O_STROCNX); a.strocny = h->plane(O_STROCNY);
        a.work1 = h->plane(W_1); a.work2 = h->plane(W_2);
        finish_kernel<<<grid2d(h), 128, 0, s>>>(a);
        ++L.n;
    }
    TRY(do_halo(L, W_1, 1, -1.0, 0.0));
    TRY(do_halo(L, W_2, 1, -1.0, 0.0));
    to_tgrid2_kernel<<<grid2d(h), 128, 0, s>>>(g, h->plane(G_TAREA), h->plane(G_UAREA), h->plane(W_1),
                                              h->plane(W_2), h->plane(O_STROCNXT), h->plane(O_STROCNYT));
    ++L.n;
    double_to_int_kernel<<<(unsigned)((h->plane_elems + 255) / 256), 256, 0, s>>>(h->plane(W_ICETMASK), h->iplane(3),
                                                                                  h->plane_elems);
    ++L.n;
    CK(cudaGetLastError());
    CK(cudaEventRecord(h->ev[3], s));
    CK(cudaMemcpyAsync(h->h_counts, h->counts(), 16, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));

    float ms;
    CK(cudaEventElapsedTime(&ms, h->ev[0], h->ev[1])); h->stats.prep_ms = ms;
    CK(cudaEventElapsedTime(&ms, h->ev[1], h->ev[2])); h->stats.loop_ms = ms;
    CK(cudaEventElapsedTime(&ms, h->ev[2], h->ev[3])); h->stats.finish_ms = ms;
    CK(cudaEventElapsedTime(&ms, h->ev[0], h->ev[3])); h->stats.total_ms = ms;
    h->stats.active_t_cells = (long long)h->h_counts[0];
    h->stats.active_u_cells = (long long)h->h_counts[1];
    h->stats.kernel_launches = L.n;
    h->stats.loop_kernel_launches = n_loop;
    return EVP_OK;
}

int evp_gpu_run(evp_handle_t *h, double dt, const evp_in_t *in, evp_inout_t *io, evp_out_t *out)
{
    if (!h || !in || !io) return EVP_ERR_ARG;
    TRY(evp_gpu_upload(h, in, io));
    TRY(evp_gpu_compute(h, dt));
    TRY(evp_gpu_download(h, io, out));
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
    b.rank = h->cfg.rank; b.device = h->cfg.device; b.pid_tag = 0;
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
