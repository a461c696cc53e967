// cuda_emu.h -- TEST INFRASTRUCTURE: a host stand-in for the slice of the CUDA runtime and of the
// device built-ins that cice5_b200/csrc uses, so that the REAL kernel and ABI sources (not a copy, not a
// port) can be compiled with g++ and run on the CPU by the `-m "not gpu"` test suite.
//
// It exists because rows of DESIGN.md section 8 were written after the round's GPU budget was spent: this is
// how their device code gets executed at all before its first hardware run.  It is NOT a product path and
// NOT a fallback: nothing under cice5_b200/ knows about it, the library it produces lives under
// tests/emu/_build/ and is loaded only by tests/emu/emu.py.  It says nothing about speed, and nothing about
// the memory model or the asynchrony of a real device: a missing fence or a race between CTAs cannot show here.
//
// Model: a kernel launch is a closure queued on its stream; a stream is a host thread that runs its queue
// in order (so several "ranks" in one process overlap exactly as on a device, and their flag spin-waits
// make progress).  Inside a launch up to 16 blocks are resident at a time and take turns (a block that
// ends makes room for the next one of the grid), so the work queue, the counters and the "last block" logic
// of a persistent grid see several CTAs in flight; the threads of a block are
// ucontext fibers scheduled round-robin, switching only at __syncthreads and at warp collectives
// (__shfl_*_sync, __ballot_sync, __activemask), which are implemented as rendezvous of the lanes named in
// the mask.  Divergence is not modelled: __activemask() is a rendezvous of the lanes of the warp that have
// not returned yet (correct for "early return, then ballot", the only way the sources use it).
// Arithmetic is plain IEEE double with -ffp-contract=off: the same operations as the strict device build.
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <ucontext.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#define EVP_EMU 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __grid_constant__
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))
#define __noinline__ __attribute__((noinline))

inline int min(int a, int b) { return a < b ? a : b; }
inline int max(int a, int b) { return a > b ? a : b; }
inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
inline long long min(long long a, long long b) { return a < b ? a : b; }
inline long long max(long long a, long long b) { return a > b ? a : b; }
inline unsigned long long min(unsigned long long a, unsigned long long b) { return a < b ? a : b; }
inline unsigned long long max(unsigned long long a, unsigned long long b) { return a > b ? a : b; }
inline double min(double a, double b) { return fmin(a, b); }
inline double max(double a, double b) { return fmax(a, b); }

struct uint3 { unsigned x, y, z; };
struct dim3 {
    unsigned x, y, z;
    dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};

namespace emu {

struct Warp {
    int lanes = 32;            // lanes that exist
    int exited = 0;
    unsigned exited_mask = 0;
    int phase = 0;             // 0 collecting, 1 releasing
    int arrived = 0, done = 0;
    unsigned arrived_mask = 0;
    uint64_t slot[32];
    unsigned pred = 0;
};

struct Block;
struct Fiber {
    ucontext_t ctx;
    uint3 tid, bid;
    dim3 bdim, gdim;
    Block *block = nullptr;
    Warp *warp = nullptr;
    int lane = 0;
    bool finished = false;
    void *stack = nullptr;
};

struct Block {
    ucontext_t sched;
    std::vector<Fiber> fibers;
    std::vector<Warp> warps;
    int nthreads = 0, exited = 0;
    int bar_phase = 0, bar_arrived = 0, bar_done = 0;
    struct Named { int phase = 0, arrived = 0, done = 0; } named[16];     // bar.sync id, n
    void *smem = nullptr;                                                  // dynamic shared memory of this block
    std::vector<std::pair<int, void *>> statics;                           // `__shared__` variables, by declaration
    int slot = 0;                                                          // which set of fiber stacks it uses
    const std::function<void()> *body = nullptr;
};

extern thread_local Fiber *cur;
void yield();                                   // back to the block scheduler
void run_grid(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body);
void *dynamic_smem();
void *block_static(int id, size_t bytes);         // storage of one `__shared__` declaration, per block
uint64_t now_ns();

// ---- warp rendezvous: every lane in `mask` deposits `v`, all read the table once it is complete --------
struct Exchange { uint64_t slot[32]; unsigned pred; unsigned mask; };
Exchange rendezvous(unsigned mask, uint64_t v, bool pred);

template <class T> inline uint64_t to_bits(T v) { uint64_t b = 0; memcpy(&b, &v, sizeof(T)); return b; }
template <class T> inline T from_bits(uint64_t b) { T v; memcpy(&v, &b, sizeof(T)); return v; }

}  // namespace emu

#define threadIdx (emu::cur->tid)
#define blockIdx (emu::cur->bid)
#define blockDim (emu::cur->bdim)
#define gridDim (emu::cur->gdim)

// ---- device built-ins ----------------------------------------------------------------------------------
void __syncthreads();
unsigned __activemask();
inline unsigned __ballot_sync(unsigned mask, int pred) { return emu::rendezvous(mask, 0, pred != 0).pred; }
template <class T> inline T __shfl_sync(unsigned mask, T v, int src)
{
    emu::Exchange e = emu::rendezvous(mask, emu::to_bits(v), false);
    return emu::from_bits<T>(e.slot[src & 31]);
}
template <class T> inline T __shfl_up_sync(unsigned mask, T v, unsigned delta)
{
    emu::Exchange e = emu::rendezvous(mask, emu::to_bits(v), false);
    const int lane = emu::cur->lane, src = lane - (int)delta;
    return src < 0 ? v : emu::from_bits<T>(e.slot[src]);
}
template <class T> inline T __shfl_down_sync(unsigned mask, T v, unsigned delta)
{
    emu::Exchange e = emu::rendezvous(mask, emu::to_bits(v), false);
    const int lane = emu::cur->lane, src = lane + (int)delta;
    return src > 31 ? v : emu::from_bits<T>(e.slot[src]);
}
inline int __popc(unsigned x) { return __builtin_popcount(x); }
template <class T> inline T __ldcg(const T *p) { return *(const volatile T *)p; }
template <class T> inline T __ldg(const T *p) { return *p; }
inline unsigned __reduce_or_sync(unsigned mask, unsigned v)
{
    emu::Exchange e = emu::rendezvous(mask, v, false);
    unsigned r = 0;
    for (int l = 0; l < 32; ++l) if (e.mask >> l & 1u) r |= (unsigned)e.slot[l];
    return r;
}
inline long long __double_as_longlong(double d) { return emu::from_bits<long long>(emu::to_bits(d)); }
inline double __longlong_as_double(long long v) { return emu::from_bits<double>(emu::to_bits(v)); }
// "shared-space addresses" are 32-bit in the PTX the sources use: a handle into a per-host-thread table of
// regions (region << 24 | offset), translated back by the mbarrier / TMA stand-ins below
size_t __cvta_generic_to_shared(const void *p);

// ---- mbarrier, bulk tensor copies, named barriers: what the TMA subcycle kernel is made of -------------
// (operands are the 32-bit shared handles above).  A copy is carried out when it is issued and then
// completes its bytes on the barrier; waiting lanes yield until the phase they wait for has flipped.
void emu_mbar_init(unsigned bar, unsigned count);
void emu_mbar_arrive(unsigned bar);
void emu_mbar_expect_tx(unsigned bar, unsigned bytes);
unsigned emu_mbar_try_wait(unsigned bar, unsigned parity);
struct CUtensorMap;
void emu_tma_load(unsigned dst, const CUtensorMap *map, unsigned bar, int rank, int x, int y, int z);
void emu_bar_sync(int id, int nthreads);
inline void __threadfence() { std::atomic_thread_fence(std::memory_order_seq_cst); }
inline void __threadfence_system() { std::atomic_thread_fence(std::memory_order_seq_cst); }
inline void __syncwarp(unsigned mask = 0xffffffffu) { emu::rendezvous(mask, 0, false); }
inline void __nanosleep(unsigned)
{
    // a spinning thread gives way to the other fibers of its block and to the other resident blocks (a CTA of the
    // persistent loop kernel waits for counters that other CTAs of the same launch publish), then to other streams
    if (emu::cur) emu::yield();
    std::this_thread::yield();
}
inline unsigned long long atomicAdd(unsigned long long *p, unsigned long long v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline unsigned atomicAdd(unsigned *p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline unsigned atomicOr(unsigned *p, unsigned v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
inline int atomicMax(int *p, int v) { int o = *p; while (o < v && !__atomic_compare_exchange_n(p, &o, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {} return o; }
inline double __dadd_rn(double a, double b) { return a + b; }
inline double __dsub_rn(double a, double b) { return a - b; }
inline double __dmul_rn(double a, double b) { return a * b; }
inline double __ddiv_rn(double a, double b) { return a / b; }
inline double __dsqrt_rn(double a) { return sqrt(a); }
inline double __fma_rn(double a, double b, double c) { return fma(a, b, c); }
// PTX of the cross-rank flags, rewritten by tests/emu/emu.py into these
inline void emu_st_release_u64(void *p, unsigned long long v) { __atomic_store_n((unsigned long long *)p, v, __ATOMIC_RELEASE); }
inline unsigned long long emu_ld_acquire_u64(const void *p) { return __atomic_load_n((const unsigned long long *)p, __ATOMIC_ACQUIRE); }
void emu_bar_arrive(int id, int nthreads);
[[noreturn]] void emu_ptx_unavailable(const char *what);

// ---- runtime API -----------------------------------------------------------------------------------------
enum cudaError_t { cudaSuccess = 0, cudaErrorInvalidValue = 1, cudaErrorMemoryAllocation = 2, cudaErrorNotSupported = 801,
                   cudaErrorPeerAccessAlreadyEnabled = 704, cudaErrorHostMemoryAlreadyRegistered = 712 };
typedef cudaError_t cudaError;
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice,
                      cudaMemcpyDefault };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16 };
enum cudaStreamCaptureMode { cudaStreamCaptureModeGlobal, cudaStreamCaptureModeThreadLocal, cudaStreamCaptureModeRelaxed };
enum cudaDriverEntryPointQueryResult { cudaDriverEntryPointSuccess = 0, cudaDriverEntryPointSymbolNotFound = 1 };
enum { cudaStreamNonBlocking = 1, cudaEnableDefault = 0, cudaIpcMemLazyEnablePeerAccess = 1 };
struct cudaDeviceProp { char name[256]; int multiProcessorCount; int major, minor; size_t totalGlobalMem; };
struct cudaIpcMemHandle_t { char reserved[64]; };
typedef cudaIpcMemHandle_t cudaIpcMemHandle;

namespace emu { struct Stream; struct Event; struct Graph; }
typedef emu::Stream *cudaStream_t;
typedef emu::Event *cudaEvent_t;
typedef emu::Graph *cudaGraph_t;
typedef emu::Graph *cudaGraphExec_t;
typedef cudaStream_t cudaStream;
typedef cudaEvent_t cudaEvent;
typedef cudaGraph_t cudaGraph;
typedef cudaGraphExec_t cudaGraphExec;

const char *cudaGetErrorString(cudaError_t);
cudaError_t cudaGetLastError();
enum { cudaHostRegisterDefault = 0 };
// page-locking: the host build keeps a set of registered ranges so that double registration is reported
cudaError_t cudaHostRegister(void *, size_t, unsigned);
cudaError_t cudaHostUnregister(void *);
cudaError_t cudaSetDevice(int);
cudaError_t cudaGetDeviceCount(int *);
cudaError_t cudaDeviceGetAttribute(int *, cudaDeviceAttr, int);
cudaError_t cudaGetDeviceProperties(cudaDeviceProp *, int);
cudaError_t cudaDeviceSynchronize();
cudaError_t cudaDeviceEnablePeerAccess(int, unsigned);
cudaError_t cudaMalloc(void **, size_t);
template <class T> inline cudaError_t cudaMalloc(T **p, size_t n) { return cudaMalloc((void **)p, n); }
cudaError_t cudaFree(void *);
cudaError_t cudaMemcpy(void *, const void *, size_t, cudaMemcpyKind);
// a `static __device__` array is a host array here: its address is the "device" address
template <class T> inline cudaError_t cudaGetSymbolAddress(void **p, const T &symbol) { *p = (void *)&symbol; return cudaSuccess; }
cudaError_t cudaMemcpyAsync(void *, const void *, size_t, cudaMemcpyKind, cudaStream_t = nullptr);
cudaError_t cudaMemcpy2DAsync(void *, size_t, const void *, size_t, size_t, size_t, cudaMemcpyKind, cudaStream_t = nullptr);
cudaError_t cudaMemset(void *, int, size_t);
cudaError_t cudaMemsetAsync(void *, int, size_t, cudaStream_t = nullptr);
cudaError_t cudaStreamCreateWithFlags(cudaStream_t *, unsigned);
cudaError_t cudaStreamDestroy(cudaStream_t);
cudaError_t cudaStreamSynchronize(cudaStream_t);
cudaError_t cudaStreamBeginCapture(cudaStream_t, cudaStreamCaptureMode);
cudaError_t cudaStreamEndCapture(cudaStream_t, cudaGraph_t *);
cudaError_t cudaGraphInstantiate(cudaGraphExec_t *, cudaGraph_t, unsigned long long);
cudaError_t cudaGraphLaunch(cudaGraphExec_t, cudaStream_t);
cudaError_t cudaGraphDestroy(cudaGraph_t);
cudaError_t cudaGraphExecDestroy(cudaGraphExec_t);
cudaError_t cudaEventCreate(cudaEvent_t *);
cudaError_t cudaEventDestroy(cudaEvent_t);
cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = nullptr);
cudaError_t cudaEventElapsedTime(float *, cudaEvent_t, cudaEvent_t);
cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0);
cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t *, void *);
cudaError_t cudaIpcOpenMemHandle(void **, cudaIpcMemHandle_t, unsigned);
cudaError_t cudaIpcCloseMemHandle(void *);
cudaError_t cudaGetDriverEntryPoint(const char *, void **, unsigned long long, cudaDriverEntryPointQueryResult * = nullptr);
template <class F> inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }

// ---- driver API types the tensor-map code names (never executed here) ------------------------------------
typedef int CUresult;
enum { CUDA_SUCCESS = 0 };
typedef uint32_t cuuint32_t;
typedef uint64_t cuuint64_t;
struct alignas(64) CUtensorMap { uint64_t opaque[16]; };
enum CUtensorMapDataType { CU_TENSOR_MAP_DATA_TYPE_UINT8 = 0, CU_TENSOR_MAP_DATA_TYPE_FLOAT64 = 10 };
enum CUtensorMapInterleave { CU_TENSOR_MAP_INTERLEAVE_NONE = 0 };
enum CUtensorMapSwizzle { CU_TENSOR_MAP_SWIZZLE_NONE = 0 };
enum CUtensorMapL2promotion { CU_TENSOR_MAP_L2_PROMOTION_NONE = 0, CU_TENSOR_MAP_L2_PROMOTION_L2_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B };
enum CUtensorMapFloatOOBfill { CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE = 0 };

// ---- launches: `k<<<grid, block, smem, stream>>>(args...)` is rewritten into emu::launch(...) ------------
namespace emu {
void enqueue(cudaStream_t s, std::function<void()> fn);
template <class F, class... A>
inline void launch(dim3 grid, dim3 block, size_t smem, cudaStream_t stream, F kernel, A... args)
{
    std::function<void()> body = [=]() { kernel(args...); };
    enqueue(stream, [=]() { run_grid(grid, block, smem, body); });
}
}  // namespace emu
