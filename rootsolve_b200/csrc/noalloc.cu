// noalloc.cu -- convergence test 1 for callers that must not allocate device memory.
//
// Who needs this: code that drives the Newton loop from inside someone else's memory regime --
// a host application that has handed the whole GPU heap to its own pool allocator, a real-time or
// in-situ coupling layer (an ocean/sediment model calling the steady-state test every time step)
// where cudaMalloc/cudaFree would serialise the device and break its latency budget, or a plugin
// running under a framework that forbids allocation once its start-up phase is over.  For those
// callers rs_convergence_test_noalloc() evaluates
//     out[0] = sum_k |beta_k|          out[1] = max_k |beta_k / ewt_k|      (src/sparse.f:122-133)
// on device buffers the caller already owns, on the caller's stream, and takes the only scratch it
// needs (per-block partial results + an arrival counter) from the fixed array below instead of
// cudaMalloc.  Nothing in this file calls cudaMalloc, cudaFree or any allocating library routine.
//
// The price of a fixed array is a fixed limit: RS_NOALLOC_SCRATCH_BYTES (64 KiB, resident on every
// GPU that loads this library until the process exits) holds partials for at most
// RS_NOALLOC_MAX_BLOCKS blocks of RS_NOALLOC_TILE unknowns, i.e. neq <= RS_NOALLOC_MAX_NEQ
// (8 355 840).  Larger systems are refused with RS_ETOOBIG; use rs_convergence_test() or the
// solver handle there.
//
// All calls on one device share the one array, and it keeps whatever the previous call left in
// it.  Each call therefore clears the arrival counter before its kernel, and two calls whose
// kernels can overlap on the GPU (different streams, no ordering between them) race on the
// partials: order them.  The first call on each device looks up the array's device address
// (cudaGetSymbolAddress) and caches it; make that first call before starting any stream capture.
#include <mutex>

#include "rs_internal.cuh"

#define RS_NOALLOC_SCRATCH_BYTES 65536
#define RS_NOALLOC_HEADER_BYTES 256            /* arrival counter, padded to the array's alignment */
#define RS_NOALLOC_TILE 2048                   /* unknowns reduced by one 256-thread block */
#define RS_NOALLOC_MAX_BLOCKS ((RS_NOALLOC_SCRATCH_BYTES - RS_NOALLOC_HEADER_BYTES) / (2 * (int)sizeof(double)))
#define RS_NOALLOC_MAX_NEQ (RS_NOALLOC_MAX_BLOCKS * RS_NOALLOC_TILE)
#define RS_NOALLOC_MAX_DEVICES 64

// In host code this name is only a stand-in for the device object: its device address comes from
// cudaGetSymbolAddress (noalloc_scratch_address below), never from taking &rs_noalloc_scratch.
static __device__ __align__(256) unsigned char rs_noalloc_scratch[RS_NOALLOC_SCRATCH_BYTES];

namespace rs {

// One block per tile of RS_NOALLOC_TILE unknowns (8 per thread, coalesced); the block results go
// to the scratch partials and the last block to arrive folds them in block order.
__global__ void __launch_bounds__(256) test1_tiled_kernel(int n, const double *__restrict__ beta,
                                                          const double *__restrict__ ewt, unsigned char *scratch,
                                                          double *__restrict__ out)
{
    unsigned *ticket = reinterpret_cast<unsigned *>(scratch);
    double *partials = reinterpret_cast<double *>(scratch + RS_NOALLOC_HEADER_BYTES);
    double acc[2] = {0.0, 0.0};
    const long long base = (long long)blockIdx.x * RS_NOALLOC_TILE;
#pragma unroll
    for (int q = 0; q < RS_NOALLOC_TILE / 256; q++) {
        const long long i = base + q * 256 + threadIdx.x;
        if (i < n) {
            const double b = beta[i];
            acc[0] += fabs(b);
            acc[1] = fmax(acc[1], fabs(b / ewt[i]));
        }
    }
    grid_reduce<2>(acc, 2u, partials, ticket, out);
}

static std::mutex g_noalloc_mutex;
static unsigned char *g_noalloc_ptr[RS_NOALLOC_MAX_DEVICES] = {nullptr};

// Device address of the scratch array on the current device; resolved once per device.
static int noalloc_scratch_address(unsigned char **ptr)
{
    int dev = 0;
    RS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= RS_NOALLOC_MAX_DEVICES) {
        set_error("rs_convergence_test_noalloc: device ordinal %d is beyond the %d this entry point tracks", dev,
                  RS_NOALLOC_MAX_DEVICES);
        return RS_ETOOBIG;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) {
        unsigned char *p = nullptr;
        RS_CUDA(cudaGetSymbolAddress((void **)&p, rs_noalloc_scratch));
        g_noalloc_ptr[dev] = p;
    }
    *ptr = g_noalloc_ptr[dev];
    return RS_OK;
}

}  // namespace rs

extern "C" int rs_convergence_test_noalloc(int neq, const double *d_beta, const double *d_ewt, double *d_out2,
                                           void *stream)
{
    using namespace rs;
    if (!d_beta || !d_ewt || !d_out2) { set_error("rs_convergence_test_noalloc: null device pointer"); return RS_EINVAL; }
    if (neq <= 0) { set_error("rs_convergence_test_noalloc: neq must be positive"); return RS_EINVAL; }
    if (neq > RS_NOALLOC_MAX_NEQ) {
        set_error("rs_convergence_test_noalloc: neq = %d exceeds the fixed-scratch limit of %d unknowns "
                  "(%d bytes of static scratch); use rs_convergence_test or a solver handle for larger systems",
                  neq, RS_NOALLOC_MAX_NEQ, RS_NOALLOC_SCRATCH_BYTES);
        return RS_ETOOBIG;
    }
    unsigned char *scratch = nullptr;
    RS_TRY(noalloc_scratch_address(&scratch));
    cudaStream_t st = (cudaStream_t)stream;
    // the array still holds the previous call's state: the arrival counter has to start at zero
    RS_CUDA(cudaMemsetAsync(scratch, 0, RS_NOALLOC_HEADER_BYTES, st));
    const int blocks = (neq + RS_NOALLOC_TILE - 1) / RS_NOALLOC_TILE;
    test1_tiled_kernel<<<blocks, 256, 0, st>>>(neq, d_beta, d_ewt, scratch, d_out2);
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

static_assert(RS_NOALLOC_MAX_NEQ == RS_CONVERGENCE_TEST_NOALLOC_MAX_NEQ,
              "the limit published in rootsolve_b200.h must match the scratch array");
