// rs_internal.cuh -- shared plumbing of the solver library: error reporting,
// device buffers, the per-call context (stream, launch counter, per-kernel
// CUDA-event statistics) and the block/grid reduction helpers.
#pragma once

#include <cuda_runtime.h>

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
    K_PRECOND_SETUP, K_PRECOND_APPLY, K_NEWTON, K_HALO, K_MISC, K_COUNT
};
extern const char *const kKernelNames[K_COUNT];

struct Ctx {
    cudaStream_t stream = nullptr;
    long long launches = 0;
    bool profile = false;  // record CUDA events around every launch
    double ms[K_COUNT] = {0};
    long long count[K_COUNT] = {0};
    double bytes[K_COUNT] = {0};  // algorithmic bytes of the last launch of the class
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
    Launch(Ctx &ctx, int kid, double algo_bytes = 0.0) : c(ctx), id(kid) { c.launches++; c.count[id]++; if (c.profile) c.begin(id, algo_bytes); else c.bytes[id] = algo_bytes; }
    ~Launch() { if (c.profile) c.end(id); }
};

// ---- device buffers ---------------------------------------------------------
extern long long g_device_bytes;  // currently allocated through DevBuf

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
        cudaError_t e = cudaMalloc((void **)&p, count * sizeof(T) + 64);
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
        if (p) { cudaFree(p); g_device_bytes -= (long long)(n * sizeof(T)); }
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

// Reduces K per-thread values over the block and, across blocks, through
// per-block partials: the last block to arrive adds the partials in block order
// (deterministic) and stores the K results in out[0..K).  is_max[k] selects max
// instead of sum.  `ticket` must be zero on entry and is reset by the last
// block.  Every thread of the block must call this.
template <int K>
__device__ __forceinline__ void grid_reduce(double (&v)[K], unsigned max_mask, double *partials,
                                            unsigned *ticket, double *out, bool accumulate = false)
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
                if (lane == 0) out[k] = accumulate ? (mx ? fmax(out[k], r) : out[k] + r) : r;
            }
            if (lane == 0) *ticket = 0;
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
