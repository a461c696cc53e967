// cuda_emu.cpp -- TEST INFRASTRUCTURE: see cuda_emu.h.
#include "cuda_emu.h"

#include <stdio.h>

#include <mutex>
#include <set>

namespace emu {

thread_local Fiber *cur = nullptr;
static thread_local Block *cur_block = nullptr;

uint64_t now_ns()
{
    return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(
               std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---- fibers ------------------------------------------------------------------------------------------
void yield()
{
    Fiber *f = cur;
    swapcontext(&f->ctx, &f->block->sched);
}

static void fiber_main()
{
    Fiber *f = cur;
    (*f->block->body)();
    f->finished = true;
    f->block->exited++;
    f->warp->exited++;
    f->warp->exited_mask |= 1u << f->lane;
    swapcontext(&f->ctx, &f->block->sched);     // never resumed
}

// stacks are kept per host thread and reused from launch to launch: one set per resident block
static const int kResident = 16;
static const size_t kStack = 96 * 1024;
static thread_local std::vector<void *> stack_pool[kResident];

// ---- 32-bit "shared" handles -----------------------------------------------------------------------------
struct Region { uintptr_t base; };
static thread_local std::vector<Region> regions;
static const uintptr_t kRegionSpan = (uintptr_t)1 << 24;

static void *from_handle(unsigned h)
{
    const unsigned r = h >> 24;
    if (r >= regions.size()) { fprintf(stderr, "cuda_emu: bad shared handle %#x\n", h); abort(); }
    return (void *)(regions[r].base + (h & 0xFFFFFFu));
}

static Block *make_block(dim3 grid, dim3 block, uint3 bid, size_t smem, const std::function<void()> &body, int slot)
{
    Block *bp = new Block;
    Block &b = *bp;
    b.nthreads = (int)(block.x * block.y * block.z);
    b.body = &body;
    b.slot = slot;
    b.fibers.resize(b.nthreads);
    b.warps.resize((b.nthreads + 31) / 32);
    std::vector<void *> &pool = stack_pool[slot];
    while ((int)pool.size() < b.nthreads) pool.push_back(malloc(kStack));
    for (int t = 0; t < b.nthreads; ++t) {
        Fiber &f = b.fibers[t];
        f.tid = uint3{(unsigned)(t % block.x), (unsigned)((t / block.x) % block.y), (unsigned)(t / (block.x * block.y))};
        f.bid = bid; f.bdim = block; f.gdim = grid;
        f.block = &b; f.warp = &b.warps[t / 32]; f.lane = t % 32;
        f.stack = pool[t];
        getcontext(&f.ctx);
        f.ctx.uc_stack.ss_sp = f.stack;
        f.ctx.uc_stack.ss_size = kStack;
        f.ctx.uc_link = nullptr;
        makecontext(&f.ctx, fiber_main, 0);
    }
    for (size_t w = 0; w < b.warps.size(); ++w) {
        const int n = b.nthreads - (int)w * 32;
        b.warps[w].lanes = n > 32 ? 32 : n;
    }
    if (smem) {
        if (posix_memalign(&b.smem, 1024, smem)) abort();
        memset(b.smem, 0xFF, smem);
        for (uintptr_t a = (uintptr_t)b.smem; a < (uintptr_t)b.smem + smem; a += kRegionSpan) regions.push_back(Region{a});
    }
    return bp;
}

static void free_block(Block *b)
{
    free(b->smem);
    for (auto &s : b->statics) free(s.second);
    delete b;
}

// one turn: every live fiber of the block runs until it yields or returns
static void turn(Block &b)
{
    for (int t = 0; t < b.nthreads; ++t) {
        Fiber &f = b.fibers[t];
        if (f.finished) continue;
        cur = &f;
        swapcontext(&b.sched, &f.ctx);
    }
}

void run_grid(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body)
{
    Fiber *saved = cur;
    regions.clear();
    const unsigned long long total = (unsigned long long)grid.x * grid.y * grid.z;
    unsigned long long next = 0;
    Block *resident[kResident] = {nullptr};
    int live = 0;
    long long idle_rounds = 0;
    auto admit = [&](int slot) {
        const unsigned x = (unsigned)(next % grid.x), y = (unsigned)((next / grid.x) % grid.y),
                       z = (unsigned)(next / ((unsigned long long)grid.x * grid.y));
        ++next;
        resident[slot] = make_block(grid, block, uint3{x, y, z}, smem, body, slot);
        ++live;
    };
    for (int s = 0; s < kResident && next < total; ++s) admit(s);
    while (live > 0) {
        bool progress = false;
        for (int s = 0; s < kResident; ++s) {
            Block *b = resident[s];
            if (!b) continue;
            const int before = b->exited;
            turn(*b);
            if (b->exited != before) progress = true;
            if (b->exited == b->nthreads) {
                free_block(b);
                resident[s] = nullptr;
                --live;
                progress = true;
                if (next < total) admit(s);
            }
        }
        if (progress) idle_rounds = 0;
        else if (++idle_rounds > 50000000LL) { fprintf(stderr, "cuda_emu: grid makes no progress (deadlocked collective?)\n"); abort(); }
    }
    cur = saved;
}

void *dynamic_smem() { return cur->block->smem; }

void *block_static(int id, size_t bytes)
{
    Block *b = cur->block;
    for (auto &s : b->statics)
        if (s.first == id) return s.second;
    void *p = nullptr;
    if (posix_memalign(&p, 64, bytes ? bytes : 1)) abort();
    memset(p, 0xFF, bytes);                      // shared memory starts undefined
    b->statics.emplace_back(id, p);
    return p;
}

// ---- collectives ---------------------------------------------------------------------------------------
Exchange rendezvous(unsigned mask, uint64_t v, bool pred)
{
    Fiber *f = cur;
    Warp *w = f->warp;
    if (w->lanes < 32) mask &= (1u << w->lanes) - 1u;
    const int need = __builtin_popcount(mask);
    while (w->phase != 0) yield();                        // the previous collective is still being read
    w->slot[f->lane] = v;
    if (pred) w->pred |= 1u << f->lane;
    w->arrived_mask |= 1u << f->lane;
    ++w->arrived;
    (void)need;
    // complete when every lane of the mask that has not returned is here (a returned lane named in a
    // full mask is tolerated, as on Volta and later)
    while (w->phase != 1) {
        if (w->arrived >= __builtin_popcount(mask & ~w->exited_mask)) { w->phase = 1; break; }
        yield();
    }
    const int readers = w->arrived;
    Exchange e;
    memcpy(e.slot, w->slot, sizeof(e.slot));
    e.pred = w->pred & mask;
    e.mask = w->arrived_mask;
    if (++w->done == readers) { w->arrived = w->done = 0; w->pred = 0; w->arrived_mask = 0; w->phase = 0; }
    return e;
}

}  // namespace emu

unsigned __activemask()
{
    // rendezvous of the lanes that have not returned: wait until everyone else has arrived here or exited
    emu::Fiber *f = emu::cur;
    emu::Warp *w = f->warp;
    while (w->phase != 0) emu::yield();
    w->arrived_mask |= 1u << f->lane;
    ++w->arrived;
    while (w->phase == 0 && w->arrived + w->exited < w->lanes) emu::yield();
    w->phase = 1;
    const unsigned m = w->arrived_mask;
    if (++w->done == w->arrived) { w->arrived = w->done = 0; w->arrived_mask = 0; w->pred = 0; w->phase = 0; }
    return m;
}

void __syncthreads()
{
    emu::Fiber *f = emu::cur;
    emu::Block *b = f->block;
    while (b->bar_phase != 0) emu::yield();
    ++b->bar_arrived;
    while (b->bar_phase == 0 && b->bar_arrived + b->exited < b->nthreads) emu::yield();
    b->bar_phase = 1;
    if (++b->bar_done == b->bar_arrived) { b->bar_arrived = b->bar_done = 0; b->bar_phase = 0; }
}

size_t __cvta_generic_to_shared(const void *p)
{
    using namespace emu;
    const uintptr_t a = (uintptr_t)p;
    for (size_t r = 0; r < regions.size(); ++r)
        if (a >= regions[r].base && a - regions[r].base < kRegionSpan) return (r << 24) | (a - regions[r].base);
    regions.push_back(Region{a & ~(uintptr_t)0xFFF});
    if (regions.size() > 255) { fprintf(stderr, "cuda_emu: too many shared regions\n"); abort(); }
    return ((regions.size() - 1) << 24) | (a - regions.back().base);
}

// ---- mbarrier: the 64-bit word in shared memory holds the whole state --------------------------------------
namespace {
struct MBar { int32_t tx; uint16_t pending; uint16_t count_phase; };   // count in bits 0-14, phase in bit 15
static_assert(sizeof(MBar) == 8, "mbarrier state is one 64-bit word");
inline void mbar_check(MBar *m)
{
    if (m->pending == 0 && m->tx == 0) {
        m->count_phase ^= 0x8000u;
        m->pending = m->count_phase & 0x7FFFu;
    }
}
}  // namespace

void emu_mbar_init(unsigned bar, unsigned count)
{
    MBar *m = (MBar *)emu::from_handle(bar);
    m->tx = 0; m->pending = (uint16_t)count; m->count_phase = (uint16_t)count;
}
void emu_mbar_arrive(unsigned bar)
{
    MBar *m = (MBar *)emu::from_handle(bar);
    if (m->pending == 0) { fprintf(stderr, "cuda_emu: mbarrier over-arrival\n"); abort(); }
    --m->pending;
    mbar_check(m);
}
void emu_mbar_expect_tx(unsigned bar, unsigned bytes)
{
    MBar *m = (MBar *)emu::from_handle(bar);
    m->tx += (int32_t)bytes;
    emu_mbar_arrive(bar);
}
unsigned emu_mbar_try_wait(unsigned bar, unsigned parity)
{
    MBar *m = (MBar *)emu::from_handle(bar);
    const unsigned phase = (m->count_phase >> 15) & 1u;
    if (phase != (parity & 1u)) return 1u;       // the phase with that parity has completed
    emu::yield();
    return 0u;
}

// ---- tensor maps: base, rank, element size, extents, byte strides, box ---------------------------------------
namespace {
struct TMap { void *base; uint64_t rank, esize, dim[3], stride[2], box[3]; };
static_assert(sizeof(TMap) <= sizeof(CUtensorMap), "tensor map fits the opaque type");
CUresult emu_encode_tiled(CUtensorMap *out, CUtensorMapDataType dt, cuuint32_t rank, void *base, const cuuint64_t *dims,
                          const cuuint64_t *strides, const cuuint32_t *box, const cuuint32_t *, CUtensorMapInterleave,
                          CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill)
{
    if (rank < 1 || rank > 3) return 1;
    TMap t;
    memset(&t, 0, sizeof(t));
    t.base = base; t.rank = rank; t.esize = dt == CU_TENSOR_MAP_DATA_TYPE_FLOAT64 ? 8 : 1;
    for (unsigned d = 0; d < 3; ++d) { t.dim[d] = d < rank ? dims[d] : 1; t.box[d] = d < rank ? box[d] : 1; }
    for (unsigned d = 0; d + 1 < rank; ++d) t.stride[d] = strides[d];
    memset(out, 0, sizeof(*out));
    memcpy(out, &t, sizeof(t));
    return CUDA_SUCCESS;
}
}  // namespace

void emu_tma_load(unsigned dst, const CUtensorMap *map, unsigned bar, int rank, int x, int y, int z)
{
    TMap t;
    memcpy(&t, map, sizeof(t));
    if ((int)t.rank != rank) { fprintf(stderr, "cuda_emu: tensor map rank mismatch\n"); abort(); }
    // the hardware faults ("illegal instruction") on a box that does not start on a 16-byte boundary of the innermost
    // dimension, or whose shared-memory destination is not 128-byte aligned
    if (((long long)x * (long long)t.esize) % 16 != 0) { fprintf(stderr, "cuda_emu: TMA box start x = %d is not 16-byte aligned\n", x); abort(); }
    char *d = (char *)emu::from_handle(dst);
    const long long c0[3] = {x, y, z};
    for (uint64_t k = 0; k < t.box[2]; ++k)
        for (uint64_t j = 0; j < t.box[1]; ++j)
            for (uint64_t i = 0; i < t.box[0]; ++i) {
                const long long c[3] = {c0[0] + (long long)i, c0[1] + (long long)j, c0[2] + (long long)k};
                const bool in = c[0] >= 0 && c[0] < (long long)t.dim[0] && c[1] >= 0 && c[1] < (long long)t.dim[1] &&
                                c[2] >= 0 && c[2] < (long long)t.dim[2];
                if (in)
                    memcpy(d, (const char *)t.base + c[0] * t.esize + c[1] * t.stride[0] + c[2] * t.stride[1], t.esize);
                else
                    memset(d, 0, t.esize);                        // out-of-bounds elements read as zero
                d += t.esize;
            }
    MBar *m = (MBar *)emu::from_handle(bar);
    m->tx -= (int32_t)(t.box[0] * t.box[1] * t.box[2] * t.esize);
    mbar_check(m);
}

void emu_bar_sync(int id, int nthreads)
{
    emu::Block *b = emu::cur->block;
    emu::Block::Named &n = b->named[id & 15];
    while (n.phase != 0) emu::yield();
    ++n.arrived;
    while (n.phase == 0 && n.arrived < nthreads) emu::yield();
    n.phase = 1;
    if (++n.done == n.arrived) { n.arrived = n.done = 0; n.phase = 0; }
}

// bar.arrive: counted like a thread that has already left (it never waits)
void emu_bar_arrive(int id, int nthreads)
{
    (void)nthreads;
    emu::Block *b = emu::cur->block;
    emu::Block::Named &n = b->named[id & 15];
    while (n.phase != 0) emu::yield();
    ++n.arrived;
    ++n.done;
}

void emu_ptx_unavailable(const char *what)
{
    fprintf(stderr, "cuda_emu: PTX that has no host stand-in was reached: %s\n", what);
    abort();
}

// ---- streams: one worker thread each, in-order queue ----------------------------------------------------
namespace emu {

struct Stream {
    std::mutex m;
    std::condition_variable cv, idle;
    std::deque<std::function<void()>> q;
    bool busy = false, stop = false, capturing = false;
    std::vector<std::function<void()>> captured;
    std::thread worker;
    Stream() : worker([this] { loop(); }) {}
    ~Stream()
    {
        { std::lock_guard<std::mutex> l(m); stop = true; }
        cv.notify_all();
        worker.join();
    }
    void loop()
    {
        for (;;) {
            std::function<void()> fn;
            {
                std::unique_lock<std::mutex> l(m);
                cv.wait(l, [this] { return stop || !q.empty(); });
                if (q.empty()) return;
                fn = std::move(q.front());
                q.pop_front();
                busy = true;
            }
            fn();
            {
                std::lock_guard<std::mutex> l(m);
                busy = false;
                if (q.empty()) idle.notify_all();
            }
        }
    }
    void push(std::function<void()> fn)
    {
        { std::lock_guard<std::mutex> l(m); q.push_back(std::move(fn)); }
        cv.notify_one();
    }
    void sync()
    {
        std::unique_lock<std::mutex> l(m);
        idle.wait(l, [this] { return q.empty() && !busy; });
    }
};

struct Event { uint64_t t = 0; std::atomic<uint64_t> issued{0}, done{0}; };
struct Graph { std::vector<std::function<void()>> nodes; };

static std::mutex g_streams_m;
static std::vector<Stream *> g_streams;
static Stream *default_stream()
{
    static Stream *s = [] { Stream *x = new Stream; std::lock_guard<std::mutex> l(g_streams_m); g_streams.push_back(x); return x; }();
    return s;
}

void enqueue(cudaStream_t s, std::function<void()> fn)
{
    if (!s) s = default_stream();
    if (s->capturing) { s->captured.push_back(std::move(fn)); return; }
    s->push(std::move(fn));
}

}  // namespace emu

// ---- runtime API -------------------------------------------------------------------------------------------
const char *cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "cuda_emu error"; }
cudaError_t cudaGetLastError() { return cudaSuccess; }
namespace { std::mutex g_reg_mu; std::set<void *> g_registered; }
cudaError_t cudaHostRegister(void *p, size_t, unsigned)
{
    std::lock_guard<std::mutex> lk(g_reg_mu);
    return g_registered.insert(p).second ? cudaSuccess : cudaErrorHostMemoryAlreadyRegistered;
}
cudaError_t cudaHostUnregister(void *p)
{
    std::lock_guard<std::mutex> lk(g_reg_mu);
    return g_registered.erase(p) ? cudaSuccess : cudaErrorInvalidValue;
}
cudaError_t cudaSetDevice(int) { return cudaSuccess; }
cudaError_t cudaGetDeviceCount(int *n) { *n = 8; return cudaSuccess; }
cudaError_t cudaDeviceGetAttribute(int *v, cudaDeviceAttr, int) { *v = 4; return cudaSuccess; }
cudaError_t cudaGetDeviceProperties(cudaDeviceProp *p, int)
{
    memset(p, 0, sizeof(*p));
    snprintf(p->name, sizeof(p->name), "cuda_emu (host)");
    p->multiProcessorCount = 4; p->major = 10; p->minor = 0; p->totalGlobalMem = (size_t)8 << 30;
    return cudaSuccess;
}
cudaError_t cudaDeviceSynchronize()
{
    std::vector<emu::Stream *> all;
    { std::lock_guard<std::mutex> l(emu::g_streams_m); all = emu::g_streams; }
    for (auto *s : all) s->sync();
    return cudaSuccess;
}
cudaError_t cudaDeviceEnablePeerAccess(int, unsigned) { return cudaSuccess; }
cudaError_t cudaMalloc(void **p, size_t n)
{
    void *q = nullptr;
    if (posix_memalign(&q, 256, n ? n : 1)) return cudaErrorMemoryAllocation;
    memset(q, 0xFF, n);                          // uninitialised device memory: NaN doubles, -1 ints
    *p = q;
    return cudaSuccess;
}
cudaError_t cudaFree(void *p) { free(p); return cudaSuccess; }
cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind)
{
    memcpy(d, s, n);                             // the streams here are all non-blocking: no implicit sync
    return cudaSuccess;
}
// Pageable host memory, as the callers of the C ABI hand in: a host-to-device copy is staged before the
// call returns (the caller may free or reuse its buffer at once), a device-to-host copy has completed when
// the call returns.
cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind k, cudaStream_t st)
{
    return cudaMemcpy2DAsync(d, n, s, n, n, 1, k, st);
}
cudaError_t cudaMemcpy2DAsync(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, cudaMemcpyKind k, cudaStream_t st)
{
    if (k == cudaMemcpyHostToDevice) {
        auto stage = std::make_shared<std::vector<char>>(w * h);
        for (size_t r = 0; r < h; ++r) memcpy(stage->data() + r * w, (const char *)s + r * sp, w);
        emu::enqueue(st, [=] { for (size_t r = 0; r < h; ++r) memcpy((char *)d + r * dp, stage->data() + r * w, w); });
        return cudaSuccess;
    }
    emu::enqueue(st, [=] { for (size_t r = 0; r < h; ++r) memcpy((char *)d + r * dp, (const char *)s + r * sp, w); });
    if (k == cudaMemcpyDeviceToHost && !(st && st->capturing)) cudaStreamSynchronize(st);
    return cudaSuccess;
}
cudaError_t cudaMemset(void *d, int v, size_t n) { memset(d, v, n); return cudaSuccess; }
cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t st)
{
    emu::enqueue(st, [=] { memset(d, v, n); });
    return cudaSuccess;
}
cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned)
{
    *s = new emu::Stream;
    std::lock_guard<std::mutex> l(emu::g_streams_m);
    emu::g_streams.push_back(*s);
    return cudaSuccess;
}
cudaError_t cudaStreamDestroy(cudaStream_t s)
{
    if (!s) return cudaSuccess;
    s->sync();
    {
        std::lock_guard<std::mutex> l(emu::g_streams_m);
        for (size_t i = 0; i < emu::g_streams.size(); ++i)
            if (emu::g_streams[i] == s) { emu::g_streams.erase(emu::g_streams.begin() + i); break; }
    }
    delete s;
    return cudaSuccess;
}
cudaError_t cudaStreamSynchronize(cudaStream_t s) { (s ? s : emu::default_stream())->sync(); return cudaSuccess; }
cudaError_t cudaStreamBeginCapture(cudaStream_t s, cudaStreamCaptureMode)
{
    s->sync();
    s->capturing = true;
    s->captured.clear();
    return cudaSuccess;
}
cudaError_t cudaStreamEndCapture(cudaStream_t s, cudaGraph_t *g)
{
    s->capturing = false;
    *g = new emu::Graph{std::move(s->captured)};
    s->captured.clear();
    return cudaSuccess;
}
cudaError_t cudaGraphInstantiate(cudaGraphExec_t *e, cudaGraph_t g, unsigned long long) { *e = new emu::Graph(*g); return cudaSuccess; }
cudaError_t cudaGraphLaunch(cudaGraphExec_t e, cudaStream_t s)
{
    for (auto &n : e->nodes) emu::enqueue(s, n);
    return cudaSuccess;
}
cudaError_t cudaGraphDestroy(cudaGraph_t g) { delete g; return cudaSuccess; }
cudaError_t cudaGraphExecDestroy(cudaGraphExec_t g) { delete g; return cudaSuccess; }
cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new emu::Event; return cudaSuccess; }
cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s)
{
    const uint64_t g = ++e->issued;
    emu::enqueue(s, [e, g] { e->t = emu::now_ns(); e->done.store(g, std::memory_order_release); });
    return cudaSuccess;
}
cudaError_t cudaStreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned)
{
    const uint64_t g = e->issued.load();         // the most recent record at the time of this call
    emu::enqueue(s, [e, g] { while (e->done.load(std::memory_order_acquire) < g) std::this_thread::yield(); });
    return cudaSuccess;
}
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b)
{
    *ms = (float)((double)(b->t - a->t) * 1e-6);
    return cudaSuccess;
}
cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t *h, void *p)
{
    memset(h, 0, sizeof(*h));
    memcpy(h->reserved, &p, sizeof(p));
    return cudaSuccess;
}
cudaError_t cudaIpcOpenMemHandle(void **p, cudaIpcMemHandle_t h, unsigned)
{
    memcpy(p, h.reserved, sizeof(*p));
    return cudaSuccess;
}
cudaError_t cudaIpcCloseMemHandle(void *) { return cudaSuccess; }
cudaError_t cudaGetDriverEntryPoint(const char *, void **fn, unsigned long long, cudaDriverEntryPointQueryResult *q)
{
    *fn = (void *)&emu_encode_tiled;             // the only driver entry point the sources ask for
    if (q) *q = cudaDriverEntryPointSuccess;
    return cudaSuccess;
}
