// rs_internal.cuh -- shared plumbing of the solver library: error reporting,
// device buffers, the per-call context (stream, launch counter, per-kernel
// CUDA-event statistics) and the block/grid reduction helpers.
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "rootsolve_b200.h"
#include "rootsolve_b200_model.cuh"

namespace rs {

void set_error(const char *fmt, ...);

#define RS_CUDA(call)                                                                          \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            rs::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return (e_ == cudaErrorMemoryAllocation) ? RS_ENOMEM : RS_ENODEVICE;               \
        }                                                                                      \
    } while (0)

#define RS_TRY(expr)                 \
    do {                             \
        int rc_ = (expr);            \
        if (rc_ != RS_OK) return rc_; \
    } while (0)

// ---- per-kernel statistics --------------------------------------------------
// Kernel classes the bench reports against the HBM roofline (DESIGN.md lists the
// algorithmic bytes of each).
enum KernelId {
    K_STRUCT = 0, K_SCAN, K_TRANSPOSE, K_JGROUP, K_PERTURB, K_RESIDUAL, K_ASSEMBLE, K_SPMV, K_VEC,
    K_PRECOND_SETUP, K_PRECOND_APPLY, K_NEWTON, K_HALO, K_MISC, K_SPMV_GRID, K_MG_SMOOTH_L0, K_MG_TRANSFER_L0, K_MG_COARSE,
    K_FD_ROWS, K_COUNT
};
extern const char *const kKernelNames[K_COUNT];

struct Ctx {
    cudaStream_t stream = nullptr;
    long long launches = 0;
    bool profile = false;  // record CUDA events around every launch
    double ms[K_COUNT] = {0};
    long long count[K_COUNT] = {0};
    double bytes[K_COUNT] = {0};  // algorithmic bytes of all launches of the class since the last reset
    struct Pending { int id; cudaEvent_t a, b; };
    std::vector<Pending> pending;
    std::vector<cudaEvent_t> pool;

    void begin(int id, double algo_bytes);
    void end(int id);
    void resolve();  // synchronises and folds pending events into ms[]
    void reset_stats();
    Ctx();           // owns a non-blocking stream (capturable into CUDA graphs, unlike the legacy default stream)
    Ctx(const Ctx &) = delete;
    Ctx &operator=(const Ctx &) = delete;
    ~Ctx();
};

// RAII guard: counts the launch and (when profiling) times it on the stream.
struct Launch {
    Ctx &c; int id;
    Launch(Ctx &ctx, int kid, double algo_bytes = 0.0) : c(ctx), id(kid) { c.launches++; c.count[id]++; if (c.profile) c.begin(id, algo_bytes); else c.bytes[id] += algo_bytes; }
    ~Launch() { if (c.profile) c.end(id); }
};

// ---- device buffers ---------------------------------------------------------
extern std::atomic<long long> g_device_bytes;  // currently allocated through DevBuf (any thread)

// Caching device allocator (api.cu).  A steady-state solve allocates ~40 arrays (gigabytes at 256^3) and frees
// them on return; cudaMalloc/cudaFree of that volume costs tens of milliseconds per call and now and then far
// more.  Freed blocks are kept per (device, size class) and handed out again; rs_release_cached_memory() gives
// them back to the driver, RS_POOL_MAX_BYTES caps what is kept (0 = no caching).
cudaError_t pool_alloc(void **p, size_t bytes);
void pool_free(void *p, size_t bytes);
// Small pinned host buffers (the scalars the host reads back): cudaMallocHost costs a millisecond or more per call --
// two of them were most of the set-up of a 50 x 50 solve -- so blocks of one fixed size are kept for reuse.
constexpr size_t PINNED_BLOCK = 512;   // bytes
cudaError_t pinned_alloc(void **p);
void pinned_free(void *p);

template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    int alloc(size_t count)
    {
        release();
        if (count == 0) count = 1;
        // 64 bytes of slack: bulk (TMA) copies round their byte ranges up to 16-byte multiples
        cudaError_t e = pool_alloc((void **)&p, count * sizeof(T) + 64);
        if (e != cudaSuccess) {
            p = nullptr;
            set_error("cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(e));
            return e == cudaErrorMemoryAllocation ? RS_ENOMEM : RS_ENODEVICE;
        }
        n = count;
        g_device_bytes += (long long)(count * sizeof(T));
        return RS_OK;
    }
    void release()
    {
        if (p) { pool_free(p, n * sizeof(T) + 64); g_device_bytes -= (long long)(n * sizeof(T)); }
        p = nullptr; n = 0;
    }
    int zero(cudaStream_t s) { RS_CUDA(cudaMemsetAsync(p, 0, n * sizeof(T), s)); return RS_OK; }
    int upload(const T *h, size_t count, cudaStream_t s) { RS_CUDA(cudaMemcpyAsync(p, h, count * sizeof(T), cudaMemcpyHostToDevice, s)); return RS_OK; }
    int download(T *h, size_t count, cudaStream_t s) const { RS_CUDA(cudaMemcpyAsync(h, p, count * sizeof(T), cudaMemcpyDeviceToHost, s)); return RS_OK; }
};

int ensure_device();  // RS_OK when a CUDA device is usable

// ---- reductions ---------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---- all-reduce over NVLink peer memory, fused into the reducing kernel ------------------------
// Every rank owns a small mailbox in device memory that all peers have mapped (CUDA IPC).  The last
// block of a reducing kernel stores its rank's partial results straight into every peer's mailbox
// (remote stores over NVLink), then waits until all peers' contributions for the same sequence
// number have landed in its own mailbox and adds them in rank order -- so every rank computes a
// bitwise identical sum, inside the kernel that produced the partial, with no extra launch.
// Slots are double-buffered by sequence parity: an all-reduce is a barrier, so a peer can run at
// most one reduction ahead.  A bounded spin turns a lost peer into NaNs (caught by the host)
// instead of a hung GPU.
constexpr int PEER_MAX = 8;
constexpr int PEER_SLOT = 4;   // doubles per (parity, source rank): two values, the sequence flag, padding
// Sequence numbers live in device memory and are advanced by the kernels themselves (only the last
// block of a launch touches them), so a launch carries no per-call argument and whole Krylov steps can
// be captured into CUDA graphs.  Every rank runs the same launches in the same order, hence the
// counters agree across ranks without ever being communicated.
struct PeerRed {
    int nranks = 1, rank = 0;
    unsigned *seq = nullptr;              // device counter: number of all-reduces completed on this rank
    double *mail[PEER_MAX] = {nullptr};   // mail[p]: rank p's mailbox as mapped into this process
};

// ---- halo push over NVLink --------------------------------------------------------------------
// The kernel that produces a vector also stores its first / last owned plane straight into the slab
// neighbours' copies (peer pointers), then the last block tells the neighbours "sequence seq is
// complete"; the consuming SpMV waits for both neighbours' flags before it touches the halo.  No NCCL
// call, no extra launch.  Flags only ever grow, so the wait is `flag >= seq`.
struct HaloSignal {
    unsigned *seq = nullptr;                               // device counter: number of halo pushes completed on this rank
    double *tell_lo = nullptr, *tell_hi = nullptr;         // flag words inside the neighbours' mailboxes
    const double *wait_lo = nullptr, *wait_hi = nullptr;   // flag words inside my own mailbox
};
struct HaloPush {
    double *lo = nullptr, *hi = nullptr;   // the neighbours' copies of the vector being written (same global indexing)
    int plane = 0;                         // unknowns per plane
    int signal = 0;                        // this launch is the last one writing the vector: raise the flags
    HaloSignal sig;
};

// store v at global index i into the neighbours' halo if i lies in the first / last plane of [b, e)
__device__ __forceinline__ void halo_store(const HaloPush &hp, long long b, long long e, long long i, double v)
{
    if (hp.lo != nullptr && i < b + hp.plane) hp.lo[i] = v;
    if (hp.hi != nullptr && i >= e - hp.plane) hp.hi[i] = v;
}

// end of a producing kernel: once every block's remote stores are out, the last block raises the flags
__device__ __forceinline__ void halo_signal_done(const HaloPush &hp, unsigned *ticket)
{
    if (!hp.signal) return;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(ticket, 1u);
        if (t == gridDim.x - 1) {
            *ticket = 0;
            const unsigned seq = *hp.sig.seq + 1u;   // this push's number; the consumer on this rank reads it back
            __threadfence_system();
            if (hp.sig.tell_lo) *(volatile double *)hp.sig.tell_lo = (double)seq;
            if (hp.sig.tell_hi) *(volatile double *)hp.sig.tell_hi = (double)seq;
            *hp.sig.seq = seq;
        }
    }
}

// start of a consuming kernel: every block waits until both neighbours have delivered the push this
// rank's own producer kernel just completed (the counter was advanced by that kernel, earlier in the stream)
__device__ __forceinline__ void halo_wait(const HaloSignal &hs)
{
    if (hs.seq == nullptr) return;
    if (threadIdx.x == 0) {
        const double want = (double)*(volatile const unsigned *)hs.seq;
        const long long t0 = clock64();
        // a neighbour that does not deliver within ~2 s is gone: stop the kernel with a launch error (the host reports it)
        // rather than compute on a stale halo
        if (hs.wait_lo) while (*(volatile const double *)hs.wait_lo < want) { if (clock64() - t0 > 4000000000LL) __trap(); }
        if (hs.wait_hi) while (*(volatile const double *)hs.wait_hi < want) { if (clock64() - t0 > 4000000000LL) __trap(); }
        __threadfence_system();
    }
    __syncthreads();
}

template <int K>
__device__ __forceinline__ void peer_allreduce_warp(double (&r)[K], const PeerRed &pr, int lane)
{
    static_assert(K <= 2, "mailbox slots carry two values");
    const unsigned seq = *(volatile const unsigned *)pr.seq + 1u;   // only the last block of a launch gets here
    const int parity = seq & 1u;
    const double flag = (double)seq;
    if (lane < pr.nranks) {
        volatile double *dst = pr.mail[lane] + (size_t)(parity * PEER_MAX + pr.rank) * PEER_SLOT;
        dst[0] = r[0];
        if (K > 1) dst[1] = r[K > 1 ? 1 : 0];
        __threadfence_system();
        dst[2] = flag;
    }
    double got[K];
#pragma unroll
    for (int k = 0; k < K; k++) got[k] = 0.0;
    if (lane < pr.nranks) {
        volatile double *src = pr.mail[pr.rank] + (size_t)(parity * PEER_MAX + lane) * PEER_SLOT;
        const long long t0 = clock64();
        bool ok = true;
        while (src[2] != flag) {
            if (clock64() - t0 > 4000000000LL) { ok = false; break; }   // ~2 s: a peer is gone
        }
        __threadfence_system();
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
        got[0] = ok ? src[0] : nan;
        if (K > 1) got[K > 1 ? 1 : 0] = ok ? src[1] : nan;
    }
    // sum in rank order on every lane -> the same bits on every rank
#pragma unroll
    for (int k = 0; k < K; k++) {
        double s = 0.0;
        for (int p = 0; p < pr.nranks; p++) s += __shfl_sync(0xffffffffu, got[k], p);
        r[k] = s;
    }
    __syncwarp();
    if (lane == 0) *pr.seq = seq;
}

// Reduces K per-thread values over the block and, across blocks, through
// per-block partials: the last block to arrive adds the partials in block order
// (deterministic) and stores the K results in out[0..K).  is_max[k] selects max
// instead of sum.  `ticket` must be zero on entry and is reset by the last
// block.  Every thread of the block must call this.  With `peer` (sum reductions
// only) the result is additionally all-reduced over the ranks before it is stored.
template <int K>
__device__ __forceinline__ void grid_reduce(double (&v)[K], unsigned max_mask, double *partials,
                                            unsigned *ticket, double *out, bool accumulate = false,
                                            const PeerRed *peer = nullptr)
{
    __shared__ double sh[K][32];
    __shared__ bool is_last;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int k = 0; k < K; k++) {
        double r = (max_mask >> k & 1u) ? warp_max(v[k]) : warp_sum(v[k]);
        if (lane == 0) sh[k][wid] = r;
    }
    __syncthreads();
    if (wid == 0) {
#pragma unroll
        for (int k = 0; k < K; k++) {
            const bool mx = max_mask >> k & 1u;
            double r = (lane < nw) ? sh[k][lane] : (mx ? -1.0e300 : 0.0);
            r = mx ? warp_max(r) : warp_sum(r);
            if (lane == 0) partials[(size_t)k * gridDim.x + blockIdx.x] = r;
        }
    }
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        if (wid == 0) {
            double res[K];
#pragma unroll
            for (int k = 0; k < K; k++) {
                const bool mx = max_mask >> k & 1u;
                // fixed order: lane l takes blocks l, l+32, ... then a shuffle tree
                double r = mx ? -1.0e300 : 0.0;
                for (unsigned b = lane; b < gridDim.x; b += 32) {
                    const double q = __ldcg(&partials[(size_t)k * gridDim.x + b]);
                    r = mx ? fmax(r, q) : r + q;
                }
                r = mx ? warp_max(r) : warp_sum(r);
                // accumulate: a launch per index range of the same reduction (stream order makes it safe)
                res[k] = accumulate ? (mx ? fmax(out[k], r) : out[k] + r) : r;
            }
            if (peer != nullptr && peer->nranks > 1) {
                if constexpr (K <= 2) peer_allreduce_warp<K>(res, *peer, lane);
            }
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < K; k++) out[k] = res[k];
                *ticket = 0;
            }
        }
    }
}

inline int grid_for(long long work, int threads, int max_blocks = 148 * 8)
{
    long long b = (work + threads - 1) / threads;
    if (b > max_blocks) b = max_blocks;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace rs
