// krylov.cu -- the Newton linear solve: right-preconditioned BiCGStab around a
// row-wise SpMV.  Stands in for the Yale sparse LDU of the reference
// (cdrv -> nroc/nsfc/nnfc/nntc, src/sparse.f:1427-2563), which re-factors the
// matrix symbolically and numerically on every Newton iteration.
//
// Kernels (all FP64, all bandwidth bound; algorithmic bytes in DESIGN.md):
//   K_SPMV   y = A x with the dot products that follow the product in BiCGStab fused
//            into the same pass.  Main variant: persistent CTAs, the matrix streamed
//            into shared memory by TMA bulk copies (one producer warp, mbarrier
//            pipeline), one consumer thread per row.  Fallback for matrices whose
//            tiles do not fit a stage: LANES threads per row + shuffle reduction.
//   K_VEC    the three fused vector updates of one BiCGStab step (direction,
//            intermediate residual, solution + residual), each also applying
//            the (block-)Jacobi preconditioner and producing its dots;
//   K_PRECOND_SETUP  inverse diagonal / inverse nspec x nspec cell blocks.
// Every scalar of the recurrence (rho, alpha, omega, beta) stays on the device:
// kernels recompute them from the raw dot products in `scal`, so one step is
// five launches and the host only reads ||r||^2 to decide when to stop.
//
// Multi-GPU: a rank works on its own index ranges (Ranges, one per species of its
// slab) of global-index vectors; every kernel takes [b, e) and a reduction that
// spans several ranges accumulates into the same slot (stream order).  The SpMV
// input gets its halo planes by ncclSend/Recv first; dots go through ncclAllReduce.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "grid_stencil.cuh"
#include "solver.cuh"

namespace rs {

constexpr int MAXS = 8;        // largest species count handled by the block preconditioner
constexpr int SLOT = 8;        // doubles per parity slot: r0v, ts, tt, rr, rhonext, bad
constexpr int S_R0V = 0, S_TS = 1, S_TT = 2, S_RR = 3, S_RHONEXT = 4, S_BAD = 5;
constexpr int S_BNORM = 2 * SLOT, S_RTRUE = 2 * SLOT + 1, S_SYM = 2 * SLOT + 2;   // S_SYM: two doubles
// slots of the CG recurrence (same parity scheme): <p,Ap>, <r,r>, <r,z> of the step, breakdown flag
constexpr int C_PQ = 0, C_RR = 1, C_RZ = 2, C_ALPHA = 3, C_BAD = 5;
struct Krylov;
static inline double own_nnz_estimate(const Krylov &k);
constexpr int NSCAL = 2 * SLOT + 8;

Krylov::~Krylov()
{
    pinned_free(h_scal);
    for (int w = 0; w < 2; w++) if (g_exec[w]) cudaGraphExecDestroy(g_exec[w]);
    if (mg) mg_destroy(mg);
    if (btd) btd_destroy(btd);
}

int pick_lanes(int n, int nnz)
{
    const double avg = n > 0 ? (double)nnz / n : 1.0;
    if (avg <= 4.0) return 4;
    if (avg <= 8.5) return 8;
    if (avg <= 16.0) return 16;
    return 32;
}

// The x gathers of the SpMV kernels go through the read-only (non-coherent) path -- except when the slab neighbours
// store the halo planes of that very vector while the kernel is already resident and waiting for them (halo push):
// read-only data must not change during the kernel's lifetime, so those launches load through L2 (ld.global.cg).
__device__ __forceinline__ double ld_x(const double *__restrict__ x, int i, bool pushed)
{
    return pushed ? __ldcg(x + i) : __ldg(x + i);
}

// ---- SpMV, sub-warp variant --------------------------------------------------------
// NDOT = 0: plain product.  NDOT = 1: out[0] = <w, y>.  NDOT = 2: out[0] = <y, w>, out[1] = <y, y>.
template <int LANES, int NDOT>
__global__ void __launch_bounds__(256) spmv_kernel(int row0, int row1, const int *__restrict__ rowptr,
                                                   const int *__restrict__ col, const double *__restrict__ val,
                                                   const double *__restrict__ x, double *__restrict__ y,
                                                   const double *__restrict__ w, double *partials, unsigned *ticket,
                                                   double *out, int accumulate, PeerRed pr, HaloSignal hw)
{
    halo_wait(hw);
    const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x % LANES;
    const long long stride = ((long long)gridDim.x * blockDim.x) / LANES;
    const long long n = row1 - row0;
    double acc[NDOT > 0 ? NDOT : 1] = {0.0};
    // every thread runs the same number of trips so the shuffles stay converged
    const long long trips = (n + stride - 1) / stride;
    long long q = gt / LANES;
    for (long long it = 0; it < trips; it++, q += stride) {
        double sum = 0.0;
        const long long r = row0 + q;
        if (q < n) {
            const int e = rowptr[r + 1];
            for (int p = rowptr[r] + lane; p < e; p += LANES) sum += val[p] * ld_x(x, col[p], hw.seq != nullptr);
        }
#pragma unroll
        for (int o = LANES / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (q < n && lane == 0) {
            y[r] = sum;
            if (NDOT == 1) acc[0] += w[r] * sum;
            if (NDOT == 2) { acc[0] += sum * w[r]; acc[1] += sum * sum; }
        }
    }
    if (NDOT > 0) grid_reduce<(NDOT > 0 ? NDOT : 1)>(acc, 0u, partials, ticket, out, accumulate != 0, &pr);
}

template <int NDOT>
static void spmv_dispatch(cudaStream_t st, int lanes, int row0, int row1, const int *rowptr, const int *col,
                          const double *val, const double *x, double *y, const double *w, double *partials,
                          unsigned *ticket, double *out, int accumulate, const PeerRed &pr, const HaloSignal &hw)
{
    const int blocks = grid_for((long long)(row1 - row0) * lanes, 256, 148 * 8);
    switch (lanes) {
    case 4: spmv_kernel<4, NDOT><<<blocks, 256, 0, st>>>(row0, row1, rowptr, col, val, x, y, w, partials, ticket, out, accumulate, pr, hw); break;
    case 8: spmv_kernel<8, NDOT><<<blocks, 256, 0, st>>>(row0, row1, rowptr, col, val, x, y, w, partials, ticket, out, accumulate, pr, hw); break;
    case 16: spmv_kernel<16, NDOT><<<blocks, 256, 0, st>>>(row0, row1, rowptr, col, val, x, y, w, partials, ticket, out, accumulate, pr, hw); break;
    default: spmv_kernel<32, NDOT><<<blocks, 256, 0, st>>>(row0, row1, rowptr, col, val, x, y, w, partials, ticket, out, accumulate, pr, hw); break;
    }
}

// algorithmic bytes of one product over `rows` rows holding `nz` nonzeros
static inline double spmv_bytes(double rows, double nz) { return 12.0 * nz + 4.0 * (rows + 1) + 16.0 * rows; }
// ... and of the kernel this solver actually runs: with implicit columns col[] (4 bytes per nonzero) is not read
static inline double spmv_bytes(const Krylov &k, double rows, double nz)
{
    return k.gridcols ? 8.0 * nz + 4.0 * (rows + 1) + 16.0 * rows : spmv_bytes(rows, nz);
}

int spmv_launch(Ctx &c, int n, int nnz, const int *rowptr, const int *col, const double *val, const double *x,
                double *y, int lanes)
{
    Launch L(c, K_SPMV, spmv_bytes(n, nnz));
    spmv_dispatch<0>(c.stream, lanes, 0, n, rowptr, col, val, x, y, nullptr, nullptr, nullptr, nullptr, 0, PeerRed(), HaloSignal());
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

// ---- SpMV, TMA-staged variant ---------------------------------------------------------
// Persistent CTAs walk tiles of TMA_ROWS consecutive rows.  The tile's slice of (val, col) is one
// contiguous byte range of each array, so an elected thread streams it into shared memory with two
// 1-D bulk copies (cp.async.bulk -> UBLKCP, completion counted on an mbarrier) TMA_STAGES tiles ahead
// of the consumers; the matrix never passes through the LSU/L1 path.  Consumers are one thread per
// row: with ascending columns the k-th entry of consecutive rows addresses consecutive x, so the x
// gathers are coalesced and mostly L1/L2 hits (the stencil bands are re-read by neighbouring rows).
//
// Tile shape from a sweep on B200 at 256^3 (scratch/tune_spmv.sh; ms per SpMV launch):
//   rows x stages: 256x4 0.384, 256x3 0.386, 128x4 0.384, 128x6 0.680, 256x6 0.682, 128x8 0.680, 512x2 0.282.
// What matters is consumer warps per SM (the x gathers are latency bound): 512 rows x 2 stages is
// 110 KB per CTA, so two CTAs = 32 consumer warps fit an SM; deeper pipelines that leave room for
// only one CTA lose.
#ifndef RS_TMA_ROWS
#define RS_TMA_ROWS 512
#endif
#ifndef RS_TMA_STAGES
#define RS_TMA_STAGES 2
#endif
#ifndef RS_TMA_CAP
#define RS_TMA_CAP (RS_TMA_ROWS * 9)
#endif
constexpr int TMA_ROWS = RS_TMA_ROWS;      // rows per tile = consumer threads per CTA
constexpr int TMA_STAGES = RS_TMA_STAGES;
constexpr int TMA_CONSUMER_WARPS = TMA_ROWS / 32;
constexpr int TMA_THREADS = TMA_ROWS + 32;   // + one producer warp
constexpr int TMA_CAP = RS_TMA_CAP;      // nonzeros a stage can hold (avg 9 per row); larger tiles -> sub-warp kernel
constexpr int TMA_STAGE_BYTES = TMA_CAP * 12;
constexpr int TMA_SMEM_BYTES = TMA_STAGES * TMA_STAGE_BYTES;

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// issues the two bulk copies of tile `tile` into a stage; a0 = first staged nonzero (multiple of 4)
__device__ __forceinline__ void tma_issue_tile(int tile, int row0, int row1, const int *__restrict__ rowptr,
                                               const int *__restrict__ col, const double *__restrict__ val,
                                               unsigned char *stage, unsigned long long *bar, int *a0_slot)
{
    const int r0 = row0 + tile * TMA_ROWS;
    const int r1 = min(row1, r0 + TMA_ROWS);
    const int p0 = rowptr[r0], p1 = rowptr[r1];
    const int a0 = p0 & ~3;
    const int cnt = ((p1 + 3) & ~3) - a0;   // multiple of 4 -> 16-byte multiples for both arrays
    *a0_slot = a0;
    if (cnt > 0) {
        mbar_expect_tx(bar, (unsigned)cnt * 12u);
        bulk_g2s(stage, val + a0, (unsigned)cnt * 8u, bar);
        bulk_g2s(stage + TMA_CAP * 8, col + a0, (unsigned)cnt * 4u, bar);
    } else {
        mbar_expect_tx(bar, 0u);
    }
}

template <int NDOT>
__global__ void __launch_bounds__(TMA_THREADS) spmv_tma_kernel(int row0, int row1, int ntiles, const int *__restrict__ rowptr,
                                                               const int *__restrict__ col, const double *__restrict__ val,
                                                               const double *__restrict__ x, double *__restrict__ y,
                                                               const double *__restrict__ w, double *partials,
                                                               unsigned *ticket, double *out, int accumulate, PeerRed pr, HaloSignal hw)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) unsigned long long full[TMA_STAGES], empty[TMA_STAGES];
    __shared__ int a0s[TMA_STAGES];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < TMA_STAGES; s++) { mbar_init(&full[s], 1u); mbar_init(&empty[s], TMA_CONSUMER_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    halo_wait(hw);   // the halo planes of x have been pushed by both neighbours
    const int my_tiles = (ntiles > (int)blockIdx.x) ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    double acc[NDOT > 0 ? NDOT : 1] = {0.0};
    if (warp == TMA_CONSUMER_WARPS) {
        // ---- producer warp: one lane keeps TMA_STAGES tiles in flight ----
        if (lane == 0) {
            for (int i = 0; i < my_tiles; i++) {
                const int s = i % TMA_STAGES;
                if (i >= TMA_STAGES) {
                    mbar_wait(&empty[s], (unsigned)(i / TMA_STAGES - 1) & 1u);   // all consumer warps left stage s
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                }
                tma_issue_tile(blockIdx.x + i * gridDim.x, row0, row1, rowptr, col, val, smem + s * TMA_STAGE_BYTES, &full[s], &a0s[s]);
            }
        }
    } else {
        // ---- consumer warps: 32 rows of every tile each, no block-wide barrier ----
        int r = row0 + (int)blockIdx.x * TMA_ROWS + tid;
        int pb = 0, pe = 0;
        if (my_tiles > 0 && r < row1) { pb = rowptr[r]; pe = rowptr[r + 1]; }
        for (int i = 0; i < my_tiles; i++) {
            const int s = i % TMA_STAGES;
            const unsigned parity = (unsigned)(i / TMA_STAGES) & 1u;
            // row pointers of the next tile are fetched while this one is processed
            const int rn = r + (int)gridDim.x * TMA_ROWS;
            int pbn = 0, pen = 0;
            if (i + 1 < my_tiles && rn < row1) { pbn = rowptr[rn]; pen = rowptr[rn + 1]; }
            mbar_wait(&full[s], parity);
            const int a0 = a0s[s];
            const double *sval = reinterpret_cast<const double *>(smem + s * TMA_STAGE_BYTES) - a0;
            const int *scol = reinterpret_cast<const int *>(smem + s * TMA_STAGE_BYTES + TMA_CAP * 8) - a0;
            // The row's (col, val) go from shared memory to registers first; the stage is handed back
            // to the producer as soon as the whole warp holds its entries, BEFORE the x gathers, so a
            // stage is busy for a TMA round trip plus a few shared-memory reads, not a gather latency.
            // Eight predicated slots per trip issue all gathers of a 5..8-entry row back to back.
            const int len = pe - pb;
            const int wtrips = __reduce_max_sync(0xffffffffu, (len + 7) >> 3);
            double sum = 0.0;
            if (wtrips == 0) {
                if (lane == 0) mbar_arrive(&empty[s]);
            }
            for (int tq = 0; tq < wtrips; tq++) {
                const int p = pb + 8 * tq;
                int c[8];
                double v[8];
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const bool ok = p + q < pe;
                    c[q] = ok ? scol[p + q] : -1;
                    v[q] = ok ? sval[p + q] : 0.0;
                }
                if (tq == wtrips - 1) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&empty[s]);   // this warp no longer reads stage s
                }
#pragma unroll
                for (int q = 0; q < 8; q++)
                    if (c[q] >= 0) sum += v[q] * ld_x(x, c[q], hw.seq != nullptr);
            }
            if (r < row1) {
                y[r] = sum;
                if (NDOT == 1) acc[0] += w[r] * sum;
                if (NDOT == 2) { acc[0] += sum * w[r]; acc[1] += sum * sum; }
            }
            r = rn; pb = pbn; pe = pen;
        }
    }
    if (NDOT > 0) grid_reduce<(NDOT > 0 ? NDOT : 1)>(acc, 0u, partials, ticket, out, accumulate != 0, &pr);
}

// ---- SpMV with implicit columns (non-cyclic regular grids) ----------------------------------------
// On a non-cyclic 1-D/2-D/3-D grid the columns of row r are a pure function of r's grid position:
// r -/+ d2*d3, r -/+ d3, r -/+ 1 where the neighbour exists, plus the same cell of the other species.
// The kernel computes them instead of streaming col[] (4 of the 12 bytes per nonzero), and -- because
// the addresses no longer depend on staged data -- issues the x gathers BEFORE it waits for the tile.
// Algorithmic bytes: 8 Z + 4 (N+1) + 16 N.  The arithmetic (entry order, one FMA chain per row) is that
// of the CSR kernels, so the result is bit-identical.  krylov_setup verifies the pattern row by row
// (grid_cols_check_kernel) before this path is taken.
template <int NSPEC>
__global__ void grid_cols_check_kernel(GridCols g, int row0, int row1, const int *__restrict__ rowptr,
                                       const int *__restrict__ col, int *__restrict__ bad)
{
    const int r = row0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= row1) return;
    GridRow<NSPEC> gr;
    gr.set(g, r, true);
    int p = rowptr[r];
    const int pe = rowptr[r + 1];
    bool ok = true;
#pragma unroll
    for (int q = 0; q < GridRow<NSPEC>::NSLOT; q++)
        if (gr.has[q]) {
            if (p >= pe || col[p] != gr.c[q]) ok = false;
            p++;
        }
    if (p != pe) ok = false;
    if (!ok) *bad = 1;
}

#ifndef RS_GRID_STAGES
#define RS_GRID_STAGES 2
#endif
#ifndef RS_GRID_MINB
#define RS_GRID_MINB 2
#endif
constexpr int GRID_STAGES = RS_GRID_STAGES;
constexpr int GRID_STAGE_BYTES = TMA_CAP * 8;
constexpr int GRID_SMEM_BYTES = GRID_STAGES * GRID_STAGE_BYTES;

template <int NDOT, int NSPEC>
__global__ void __launch_bounds__(TMA_THREADS, RS_GRID_MINB)
spmv_grid_kernel(GridCols g, int row0, int row1, int ntiles, const int *__restrict__ rowptr, const double *__restrict__ val,
                 const double *__restrict__ x, double *__restrict__ y, const double *__restrict__ w, double *partials,
                 unsigned *ticket, double *out, int accumulate, PeerRed pr, HaloSignal hw)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) unsigned long long full[GRID_STAGES], empty[GRID_STAGES];
    __shared__ int a0s[GRID_STAGES];
    constexpr int NSLOT = GridRow<NSPEC>::NSLOT;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < GRID_STAGES; s++) { mbar_init(&full[s], 1u); mbar_init(&empty[s], TMA_CONSUMER_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    halo_wait(hw);
    const int my_tiles = (ntiles > (int)blockIdx.x) ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    double acc[NDOT > 0 ? NDOT : 1] = {0.0};
    if (warp == TMA_CONSUMER_WARPS) {
        if (lane == 0) {
            for (int i = 0; i < my_tiles; i++) {
                const int s = i % GRID_STAGES;
                if (i >= GRID_STAGES) {
                    mbar_wait(&empty[s], (unsigned)(i / GRID_STAGES - 1) & 1u);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                }
                const int r0 = row0 + ((int)blockIdx.x + i * (int)gridDim.x) * TMA_ROWS;
                const int r1 = min(row1, r0 + TMA_ROWS);
                const int p0 = rowptr[r0], p1 = rowptr[r1];
                const int a0 = p0 & ~1;                      // 16-byte aligned start of the value slice
                const int cnt = ((p1 + 1) & ~1) - a0;
                a0s[s] = a0;
                mbar_expect_tx(&full[s], (unsigned)cnt * 8u);
                if (cnt > 0) bulk_g2s(smem + s * GRID_STAGE_BYTES, val + a0, (unsigned)cnt * 8u, &full[s]);
            }
        }
    } else {
        int r = row0 + (int)blockIdx.x * TMA_ROWS + tid;
        int pb = (my_tiles > 0 && r < row1) ? rowptr[r] : 0;
        for (int i = 0; i < my_tiles; i++) {
            const int s = i % GRID_STAGES;
            const unsigned parity = (unsigned)(i / GRID_STAGES) & 1u;
            const int rn = r + (int)gridDim.x * TMA_ROWS;
            const int pbn = (i + 1 < my_tiles && rn < row1) ? rowptr[rn] : 0;
            // the gathers need nothing from the staged tile: they are in flight while the tile arrives
            GridRow<NSPEC> gr;
            gr.set(g, r, r < row1);
            double xv[NSLOT];
#pragma unroll
            for (int q = 0; q < NSLOT; q++) xv[q] = gr.has[q] ? ld_x(x, gr.c[q], hw.seq != nullptr) : 0.0;
            mbar_wait(&full[s], parity);
            const double *sval = reinterpret_cast<const double *>(smem + s * GRID_STAGE_BYTES) - a0s[s];
            double v[NSLOT];
            int p = pb;
#pragma unroll
            for (int q = 0; q < NSLOT; q++) {
                v[q] = gr.has[q] ? sval[p] : 0.0;
                p += gr.has[q] ? 1 : 0;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);   // the stage goes back before the gathers are waited for
            double sum = 0.0;
#pragma unroll
            for (int q = 0; q < NSLOT; q++)
                if (gr.has[q]) sum += v[q] * xv[q];
            if (r < row1) {
                y[r] = sum;
                if (NDOT == 1) acc[0] += w[r] * sum;
                if (NDOT == 2) { acc[0] += sum * w[r]; acc[1] += sum * sum; }
            }
            r = rn; pb = pbn;
        }
    }
    if (NDOT > 0) grid_reduce<(NDOT > 0 ? NDOT : 1)>(acc, 0u, partials, ticket, out, accumulate != 0, &pr);
}

// largest number of staged nonzeros any tile of [row0, row1) needs
__global__ void tma_tile_max_kernel(int row0, int row1, int ntiles, const int *__restrict__ rowptr, int *__restrict__ out)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    const int r0 = row0 + t * TMA_ROWS, r1 = min(row1, r0 + TMA_ROWS);
    const int need = ((rowptr[r1] + 3) & ~3) - (rowptr[r0] & ~3);
    atomicMax(out, need);
}

// y = A x over the rank's rows, with NDOT fused dots (accumulated over the ranges into out[])
template <int NDOT>
static void spmv_run(const Krylov &k, cudaStream_t st, const double *x, double *y, const double *w, double *out,
                     const PeerRed &peer, const HaloSignal &hw)
{
    for (int q = 0; q < k.own.n; q++) {
        const int row0 = k.own.b[q], row1 = k.own.e[q];
        if (row1 <= row0) continue;
        const PeerRed pr = (q == k.own.n - 1) ? peer : PeerRed();   // the ranks meet once, after the last range
        if (k.gridcols) {
            const int ntiles = (row1 - row0 + TMA_ROWS - 1) / TMA_ROWS;
            const int grid = std::min(ntiles, k.tma_sms * k.tma_per_sm_grid);
            const GridCols g{k.gd1, k.gd2, k.gd3, k.gd2 * k.gd3, k.ncell};
            if (k.gridcols == 1)
                spmv_grid_kernel<NDOT, 1><<<grid, TMA_THREADS, GRID_SMEM_BYTES, st>>>(g, row0, row1, ntiles, k.rowptr, k.val, x, y, w,
                                                                                     k.partials.p, k.ticket.p, out, q > 0, pr, hw);
            else
                spmv_grid_kernel<NDOT, 2><<<grid, TMA_THREADS, GRID_SMEM_BYTES, st>>>(g, row0, row1, ntiles, k.rowptr, k.val, x, y, w,
                                                                                     k.partials.p, k.ticket.p, out, q > 0, pr, hw);
        } else if (k.tma_ok) {
            const int ntiles = (row1 - row0 + TMA_ROWS - 1) / TMA_ROWS;
            const int grid = std::min(ntiles, k.tma_sms * k.tma_per_sm);
            spmv_tma_kernel<NDOT><<<grid, TMA_THREADS, TMA_SMEM_BYTES, st>>>(row0, row1, ntiles, k.rowptr, k.col, k.val, x, y, w,
                                                                             k.partials.p, k.ticket.p, out, q > 0, pr, hw);
        } else {
            spmv_dispatch<NDOT>(st, k.lanes, row0, row1, k.rowptr, k.col, k.val, x, y, w, k.partials.p, k.ticket.p, out, q > 0, pr, hw);
        }
    }
}

// ---- preconditioner set-up ----------------------------------------------------------
__device__ __forceinline__ double csr_entry(const int *__restrict__ rowptr, const int *__restrict__ col,
                                            const double *__restrict__ val, int i, int j)
{
    int b = rowptr[i], e = rowptr[i + 1];
    while (b < e) {
        const int mid = (b + e) >> 1;
        const int cj = col[mid];
        if (cj == j) return val[mid];
        if (cj < j) b = mid + 1; else e = mid;
    }
    return 0.0;
}

__global__ void jacobi_setup_kernel(int row0, int row1, const int *__restrict__ rowptr, const int *__restrict__ col,
                                    const double *__restrict__ val, double *__restrict__ dinv)
{
    const int i = row0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row1) return;
    const double d = csr_entry(rowptr, col, val, i, i);
    dinv[i] = (d != 0.0) ? 1.0 / d : 1.0;
}

// One thread per cell: gathers the nspec x nspec block coupling the species of the
// cell (the "other species in the same box" entries of src/sparsecoeff.f:70-75,
// 162-166, 296-302 plus the diagonal) and inverts it by Gauss-Jordan with
// partial pivoting.  binv is stored [s][s'][cell] so that applying it is coalesced.
__global__ void block_setup_kernel(int nspec, int ncell, int cell0, int cell1, const int *__restrict__ rowptr,
                                   const int *__restrict__ col, const double *__restrict__ val, double *__restrict__ binv)
{
    const int cidx = cell0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (cidx >= cell1) return;
    double A[MAXS][MAXS], B[MAXS][MAXS];
    for (int s = 0; s < nspec; s++)
        for (int q = 0; q < nspec; q++) {
            A[s][q] = csr_entry(rowptr, col, val, s * ncell + cidx, q * ncell + cidx);
            B[s][q] = (s == q) ? 1.0 : 0.0;
        }
    for (int k = 0; k < nspec; k++) {
        int piv = k;
        double best = fabs(A[k][k]);
        for (int s = k + 1; s < nspec; s++)
            if (fabs(A[s][k]) > best) { best = fabs(A[s][k]); piv = s; }
        if (best == 0.0) {  // singular block: fall back to the identity for this cell
            for (int s = 0; s < nspec; s++)
                for (int q = 0; q < nspec; q++) B[s][q] = (s == q) ? 1.0 : 0.0;
            break;
        }
        if (piv != k)
            for (int q = 0; q < nspec; q++) {
                double tA = A[k][q]; A[k][q] = A[piv][q]; A[piv][q] = tA;
                double tB = B[k][q]; B[k][q] = B[piv][q]; B[piv][q] = tB;
            }
        const double inv = 1.0 / A[k][k];
        for (int q = 0; q < nspec; q++) { A[k][q] *= inv; B[k][q] *= inv; }
        for (int s = 0; s < nspec; s++) {
            if (s == k) continue;
            const double f = A[s][k];
            if (f == 0.0) continue;
            for (int q = 0; q < nspec; q++) { A[s][q] -= f * A[k][q]; B[s][q] -= f * B[k][q]; }
        }
    }
    for (int s = 0; s < nspec; s++)
        for (int q = 0; q < nspec; q++) binv[((long long)s * nspec + q) * ncell + cidx] = B[s][q];
}

// ---- fused vector updates --------------------------------------------------------------
// Element loops run over an index range [b, e) with 16-byte accesses: the even-aligned interior goes
// two unknowns per thread and trip (every vector base is 256-byte aligned, so index parity decides
// alignment); a possible odd first / last element is handled by one thread.
#define RS_RANGE_LOOP(b, e, ONE, TWO)                                                                  \
    do {                                                                                               \
        const long long b_al_ = ((long long)(b) + 1) & ~1LL, e_al_ = (long long)(e) & ~1LL;            \
        const long long stride_ = (long long)gridDim.x * blockDim.x;                                   \
        if (b_al_ < e_al_) {                                                                           \
            _Pragma("unroll 2")                                                                        \
            for (long long i2 = (b_al_ >> 1) + (long long)blockIdx.x * blockDim.x + threadIdx.x;       \
                 i2 < (e_al_ >> 1); i2 += stride_) { TWO(i2); }                                        \
            if (blockIdx.x == 0 && threadIdx.x == 0) {                                                 \
                if ((b) < b_al_) { ONE((long long)(b)); }                                              \
                if (e_al_ < (e)) { ONE(e_al_); }                                                       \
            }                                                                                          \
        } else if (blockIdx.x == 0 && threadIdx.x == 0) {                                              \
            for (long long i1 = (b); i1 < (e); i1++) { ONE(i1); }                                      \
        }                                                                                              \
    } while (0)

// PRE: 0 none, 1 Jacobi, 2 block-Jacobi (work item = one cell, all its species).

// p = r + beta (p - omega v)  [first step: p = r] ; ph = M^-1 p
template <int PRE>
__global__ void __launch_bounds__(256) update_p_kernel(int b, int e, int nspec, int ncell, int first, const double *scal,
                                                       int cur, const double *__restrict__ r, const double *__restrict__ v,
                                                       double *__restrict__ p, double *__restrict__ ph,
                                                       const double *__restrict__ dinv, const double *__restrict__ binv,
                                                       HaloPush hp, unsigned *hticket)
{
    double beta = 0.0, omega = 0.0;
    if (!first) {
        const double *prev = scal + SLOT * (cur ^ 1), *curp = scal + SLOT * cur;
        const double rho_k = prev[S_RHONEXT], rho_km1 = curp[S_RHONEXT];
        const double alpha = rho_km1 / prev[S_R0V];
        omega = prev[S_TS] / prev[S_TT];
        beta = (rho_k / rho_km1) * (alpha / omega);
        if (!isfinite(beta)) { beta = 0.0; omega = 0.0; }
    }
    if (PRE == 2) {
        const long long stride = (long long)gridDim.x * blockDim.x;
        for (long long c = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; c < e; c += stride) {
            double pv[MAXS];
            for (int s = 0; s < nspec; s++) {
                const long long i = (long long)s * ncell + c;
                const double pn = first ? r[i] : r[i] + beta * (p[i] - omega * v[i]);
                pv[s] = pn;
                p[i] = pn;
            }
            for (int s = 0; s < nspec; s++) {
                double a = 0.0;
                for (int q = 0; q < nspec; q++) a += binv[((long long)s * nspec + q) * ncell + c] * pv[q];
                ph[(long long)s * ncell + c] = a;
                if (hp.lo != nullptr && c < b + hp.plane) hp.lo[(long long)s * ncell + c] = a;
                if (hp.hi != nullptr && c >= e - hp.plane) hp.hi[(long long)s * ncell + c] = a;
            }
        }
    } else {
        const double2 *r2 = reinterpret_cast<const double2 *>(r), *v2 = reinterpret_cast<const double2 *>(v);
        const double2 *d2 = reinterpret_cast<const double2 *>(dinv);
        double2 *p2 = reinterpret_cast<double2 *>(p), *ph2 = reinterpret_cast<double2 *>(ph);
#define P_ONE(i) { const double pn = first ? r[i] : r[i] + beta * (p[i] - omega * v[i]); p[i] = pn; const double ph_ = (PRE == 1) ? dinv[i] * pn : pn; \
                   ph[i] = ph_; halo_store(hp, b, e, i, ph_); }
#define P_TWO(i) { double2 pn = r2[i]; if (!first) { const double2 po = p2[i], vo = v2[i]; pn.x += beta * (po.x - omega * vo.x); pn.y += beta * (po.y - omega * vo.y); } \
                   p2[i] = pn; if (PRE == 1) { const double2 d = d2[i]; pn.x *= d.x; pn.y *= d.y; } ph2[i] = pn; \
                   halo_store(hp, b, e, 2 * (i), pn.x); halo_store(hp, b, e, 2 * (i) + 1, pn.y); }
        RS_RANGE_LOOP(b, e, P_ONE, P_TWO);
#undef P_ONE
#undef P_TWO
    }
    halo_signal_done(hp, hticket);
}

// s = r - alpha v ; sh = M^-1 s
template <int PRE>
__global__ void __launch_bounds__(256) update_s_kernel(int b, int e, int nspec, int ncell, const double *scal, int cur,
                                                       const double *__restrict__ r, const double *__restrict__ v,
                                                       double *__restrict__ s_, double *__restrict__ sh,
                                                       const double *__restrict__ dinv, const double *__restrict__ binv,
                                                       HaloPush hp, unsigned *hticket)
{
    const double rho = scal[SLOT * (cur ^ 1) + S_RHONEXT];
    double alpha = rho / scal[SLOT * cur + S_R0V];
    if (!isfinite(alpha)) alpha = 0.0;  // breakdown: flagged by update_xr_kernel, host restarts
    if (PRE == 2) {
        const long long stride = (long long)gridDim.x * blockDim.x;
        for (long long c = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; c < e; c += stride) {
            double sv[MAXS];
            for (int s = 0; s < nspec; s++) {
                const long long i = (long long)s * ncell + c;
                sv[s] = r[i] - alpha * v[i];
                s_[i] = sv[s];
            }
            for (int s = 0; s < nspec; s++) {
                double a = 0.0;
                for (int q = 0; q < nspec; q++) a += binv[((long long)s * nspec + q) * ncell + c] * sv[q];
                sh[(long long)s * ncell + c] = a;
                if (hp.lo != nullptr && c < b + hp.plane) hp.lo[(long long)s * ncell + c] = a;
                if (hp.hi != nullptr && c >= e - hp.plane) hp.hi[(long long)s * ncell + c] = a;
            }
        }
    } else {
        const double2 *r2 = reinterpret_cast<const double2 *>(r), *v2 = reinterpret_cast<const double2 *>(v);
        const double2 *d2 = reinterpret_cast<const double2 *>(dinv);
        double2 *s2 = reinterpret_cast<double2 *>(s_), *sh2 = reinterpret_cast<double2 *>(sh);
#define S_ONE(i) { const double sn = r[i] - alpha * v[i]; s_[i] = sn; const double sh_ = (PRE == 1) ? dinv[i] * sn : sn; sh[i] = sh_; halo_store(hp, b, e, i, sh_); }
#define S_TWO(i) { const double2 ro = r2[i], vo = v2[i]; double2 sn; sn.x = ro.x - alpha * vo.x; sn.y = ro.y - alpha * vo.y; s2[i] = sn; \
                   if (PRE == 1) { const double2 d = d2[i]; sn.x *= d.x; sn.y *= d.y; } sh2[i] = sn; \
                   halo_store(hp, b, e, 2 * (i), sn.x); halo_store(hp, b, e, 2 * (i) + 1, sn.y); }
        RS_RANGE_LOOP(b, e, S_ONE, S_TWO);
#undef S_ONE
#undef S_TWO
    }
    halo_signal_done(hp, hticket);
}

// x += alpha ph + omega sh ; r = s - omega t ; out = { <r,r>, <r0,r> }
__global__ void __launch_bounds__(256) update_xr_kernel(int b, int e, double *scal, int cur, const double *__restrict__ ph,
                                                        const double *__restrict__ sh, const double *__restrict__ s_,
                                                        const double *__restrict__ t, const double *__restrict__ r0,
                                                        double *__restrict__ x, double *__restrict__ r,
                                                        double *partials, unsigned *ticket, int accumulate, PeerRed pr)
{
    double *slot = scal + SLOT * cur;
    const double rho = scal[SLOT * (cur ^ 1) + S_RHONEXT];
    double alpha = rho / slot[S_R0V];
    double omega = slot[S_TS] / slot[S_TT];
    // A zero or non-finite recurrence scalar is a BiCGStab breakdown.  Leave x untouched (alpha =
    // omega = 0 keeps x and turns r into s = r) and raise the flag the host restarts on.
    // s = 0 to the last bit (t = A M^-1 s = 0, <t,t> = 0): the half step x + alpha ph solved the system -- what an exact
    // preconditioner (the block-tridiagonal solve on a small 1-D grid) produces.  Not a breakdown: take the half step.
    const bool lucky = isfinite(alpha) && rho != 0.0 && slot[S_TT] == 0.0;
    if (lucky) omega = 0.0;
    const bool bad = !lucky && (!(isfinite(alpha) && isfinite(omega)) || rho == 0.0 || omega == 0.0);
    if (bad) { alpha = 0.0; omega = 0.0; }
    if (blockIdx.x == 0 && threadIdx.x == 0) slot[S_BAD] = bad ? 1.0 : 0.0;
    double acc[2] = {0.0, 0.0};
    const double2 *ph2 = reinterpret_cast<const double2 *>(ph), *sh2 = reinterpret_cast<const double2 *>(sh);
    const double2 *s2 = reinterpret_cast<const double2 *>(s_), *t2 = reinterpret_cast<const double2 *>(t);
    const double2 *r02 = reinterpret_cast<const double2 *>(r0);
    double2 *x2 = reinterpret_cast<double2 *>(x), *r2 = reinterpret_cast<double2 *>(r);
#define X_ONE(i) { x[i] += alpha * ph[i] + omega * sh[i]; const double rn = s_[i] - omega * t[i]; r[i] = rn; acc[0] += rn * rn; acc[1] += r0[i] * rn; }
#define X_TWO(i) { const double2 a = ph2[i], bb = sh2[i], sv = s2[i], tv = t2[i], q = r02[i]; double2 xv = x2[i], rn; \
                   xv.x += alpha * a.x + omega * bb.x; xv.y += alpha * a.y + omega * bb.y; x2[i] = xv; \
                   rn.x = sv.x - omega * tv.x; rn.y = sv.y - omega * tv.y; r2[i] = rn; \
                   acc[0] += rn.x * rn.x + rn.y * rn.y; acc[1] += q.x * rn.x + q.y * rn.y; }
    RS_RANGE_LOOP(b, e, X_ONE, X_TWO);
#undef X_ONE
#undef X_TWO
    grid_reduce<2>(acc, 0u, partials, ticket, slot + S_RR, accumulate != 0, &pr);
}

// r = r0 = b ; x = 0 ; out = <b,b>
__global__ void __launch_bounds__(256) krylov_init_kernel(int b, int e, const double *__restrict__ rhs, double *__restrict__ r,
                                                          double *__restrict__ r0, double *__restrict__ x,
                                                          double *partials, unsigned *ticket, double *scal, int accumulate, PeerRed pr)
{
    double acc[1] = {0.0};
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) {
        const double bi = rhs[i];
        r[i] = bi; r0[i] = bi; x[i] = 0.0;
        acc[0] += bi * bi;
    }
    grid_reduce<1>(acc, 0u, partials, ticket, scal + S_BNORM, accumulate != 0, &pr);
}

// restart: r = r0 = b - Ax (Ax given) ; out = <r,r> (the true residual)
__global__ void __launch_bounds__(256) krylov_restart_kernel(int b, int e, const double *__restrict__ rhs,
                                                             const double *__restrict__ ax, double *__restrict__ r,
                                                             double *__restrict__ r0, double *partials,
                                                             unsigned *ticket, double *scal, int accumulate, PeerRed pr)
{
    double acc[1] = {0.0};
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) {
        const double ri = rhs[i] - ax[i];
        r[i] = ri; r0[i] = ri;
        acc[0] += ri * ri;
    }
    grid_reduce<1>(acc, 0u, partials, ticket, scal + S_RTRUE, accumulate != 0, &pr);
}

// rho_0 = <r0, r> = ||r||^2 goes where step 0 looks for "rho of the previous step"
__global__ void krylov_seed_kernel(double *scal, int src)
{
    scal[SLOT * 1 + S_RHONEXT] = scal[src];
}

// ---- host driver ------------------------------------------------------------------------
int krylov_setup(Ctx &c, Krylov &k, int n, int nnz, const int *rowptr, const int *col, const double *val,
                 int precond, int nspec, int ncell, const rs_krylov_control *ctl, const Ranges *own,
                 const HaloPlan *halo, const GridDesc *grid)
{
    k.n = n; k.nnz = nnz; k.rowptr = rowptr; k.col = col; k.val = val;
    k.nspec = nspec; k.ncell = ncell;
    if (own) k.own = *own; else { k.own.n = 1; k.own.b[0] = 0; k.own.e[0] = n; }
    if (halo) k.halo = *halo; else k.halo = HaloPlan();
    k.cell0 = 0; k.cell1 = ncell;
    if (halo && halo->nspec > 0) { k.cell0 = halo->j0 * halo->plane; k.cell1 = halo->j1 * halo->plane; }
    // multigrid on 2-D / 3-D grids (asked for, or AUTO); the (block-)Jacobi preconditioner below stays the base: the
    // fallback should a solve with the cycle fail, and what every other structure uses
    const bool want_mg = (precond == RS_PRECOND_MG || precond == RS_PRECOND_AUTO) && grid && !getenv("RS_NO_MG") &&
                         mg_applicable(*grid, k.halo.nspec > 0 ? k.halo.j0 : 0, k.halo.nspec > 0 ? k.halo.j1 : 0);
    // block-tridiagonal solve: 1-D grids (blocks of nspec), general structures (scalar band; AUTO decides at the first
    // update whether the band carries enough of the weight)
    const bool is_1d = grid && grid->type == RS_TYPE_1D && nspec <= MAXS;
    const bool want_btd = !want_mg && !getenv("RS_NO_BTD") && n >= 2 &&
                          ((precond == RS_PRECOND_BTD && (is_1d || !grid || !grid->type)) || (precond == RS_PRECOND_AUTO && (is_1d || !grid || !grid->type)));
    k.btd_decide = (want_btd && !is_1d && precond == RS_PRECOND_AUTO) ? 1 : 0;
    if (precond == RS_PRECOND_MG || precond == RS_PRECOND_BTD) precond = RS_PRECOND_AUTO;
    if (precond == RS_PRECOND_AUTO) precond = (nspec > 1 && nspec <= MAXS && ncell > 0) ? RS_PRECOND_BLOCK_JACOBI : RS_PRECOND_JACOBI;
    if (precond == RS_PRECOND_BLOCK_JACOBI && !(nspec > 1 && nspec <= MAXS && ncell > 0 && (long long)nspec * ncell == n))
        precond = RS_PRECOND_JACOBI;
    k.precond = precond;
    if (ctl) {
        if (ctl->max_krylov > 0) k.max_iter = ctl->max_krylov;
        if (ctl->lin_rtol > 0) k.lin_rtol = ctl->lin_rtol;
        if (ctl->lin_ewt_factor > 0) k.lin_ewt_factor = ctl->lin_ewt_factor;
        if (ctl->check_every > 0) k.check_every = ctl->check_every;
    }
    k.lanes = pick_lanes(n, nnz);
    // TMA-staged SpMV when every tile's slice fits one stage (RS_SPMV=subwarp forces the other kernel)
    k.tma_ok = 0;
    {
        const char *force = getenv("RS_SPMV");
        if (!(force && strcmp(force, "subwarp") == 0)) {
            DevBuf<int> mx;
            RS_TRY(mx.alloc(1));
            RS_TRY(mx.zero(c.stream));
            for (int q = 0; q < k.own.n; q++) {
                const int rows = k.own.e[q] - k.own.b[q];
                if (rows <= 0) continue;
                const int ntiles = (rows + TMA_ROWS - 1) / TMA_ROWS;
                tma_tile_max_kernel<<<(ntiles + 255) / 256, 256, 0, c.stream>>>(k.own.b[q], k.own.e[q], ntiles, rowptr, mx.p);
                c.launches++;
            }
            int need = 0;
            RS_TRY(mx.download(&need, 1, c.stream));
            RS_CUDA(cudaStreamSynchronize(c.stream));
            if (need <= TMA_CAP) {
                RS_CUDA(cudaFuncSetAttribute(spmv_tma_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, TMA_SMEM_BYTES));
                RS_CUDA(cudaFuncSetAttribute(spmv_tma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TMA_SMEM_BYTES));
                RS_CUDA(cudaFuncSetAttribute(spmv_tma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TMA_SMEM_BYTES));
                int dev = 0;
                RS_CUDA(cudaGetDevice(&dev));
                RS_CUDA(cudaDeviceGetAttribute(&k.tma_sms, cudaDevAttrMultiProcessorCount, dev));
                RS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&k.tma_per_sm, spmv_tma_kernel<2>, TMA_THREADS, TMA_SMEM_BYTES));
                if (k.tma_per_sm >= 1) k.tma_ok = 1;
            }
        }
    }
    // implicit columns: non-cyclic grid, 1 or 2 species, pattern verified row by row against the stencil
    k.gridcols = 0;
    if (k.tma_ok && grid && grid->type && !getenv("RS_NO_GRIDCOLS") && (grid->nspec == 1 || grid->nspec == 2) &&
        !grid->cyc[0] && !grid->cyc[1] && !grid->cyc[2] && (long long)grid->nspec * grid->ncell == n) {
        const int nd = grid->type == RS_TYPE_1D ? 1 : grid->type == RS_TYPE_2D ? 2 : 3;
        k.gd1 = nd == 3 ? grid->d[0] : 1;
        k.gd2 = nd == 3 ? grid->d[1] : nd == 2 ? grid->d[0] : 1;
        k.gd3 = grid->d[nd - 1];
        const GridCols g{k.gd1, k.gd2, k.gd3, k.gd2 * k.gd3, grid->ncell};
        DevBuf<int> bad;
        RS_TRY(bad.alloc(1));
        RS_TRY(bad.zero(c.stream));
        for (int q = 0; q < k.own.n; q++) {
            const int rows = k.own.e[q] - k.own.b[q];
            if (rows <= 0) continue;
            Launch L(c, K_PRECOND_SETUP, 4.0 * nnz * rows / n + 8.0 * rows);
            if (grid->nspec == 1) grid_cols_check_kernel<1><<<(rows + 255) / 256, 256, 0, c.stream>>>(g, k.own.b[q], k.own.e[q], rowptr, col, bad.p);
            else grid_cols_check_kernel<2><<<(rows + 255) / 256, 256, 0, c.stream>>>(g, k.own.b[q], k.own.e[q], rowptr, col, bad.p);
        }
        int hbad = 1;
        RS_TRY(bad.download(&hbad, 1, c.stream));
        RS_CUDA(cudaStreamSynchronize(c.stream));
        if (!hbad) {
            int per_sm = 0;
            if (grid->nspec == 1) {
                RS_CUDA(cudaFuncSetAttribute(spmv_grid_kernel<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_SMEM_BYTES));
                RS_CUDA(cudaFuncSetAttribute(spmv_grid_kernel<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_SMEM_BYTES));
                RS_CUDA(cudaFuncSetAttribute(spmv_grid_kernel<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_SMEM_BYTES));
                RS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spmv_grid_kernel<2, 1>, TMA_THREADS, GRID_SMEM_BYTES));
            } else {
                RS_CUDA(cudaFuncSetAttribute(spmv_grid_kernel<0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_SMEM_BYTES));
                RS_CUDA(cudaFuncSetAttribute(spmv_grid_kernel<1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_SMEM_BYTES));
                RS_CUDA(cudaFuncSetAttribute(spmv_grid_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_SMEM_BYTES));
                RS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spmv_grid_kernel<2, 2>, TMA_THREADS, GRID_SMEM_BYTES));
            }
            if (per_sm >= 1) { k.tma_per_sm_grid = per_sm; k.gridcols = grid->nspec; }
        }
    }
    RS_TRY(k.r.alloc(n)); RS_TRY(k.r0.alloc(n)); RS_TRY(k.p.alloc(n)); RS_TRY(k.v.alloc(n));
    RS_TRY(k.s.alloc(n)); RS_TRY(k.t.alloc(n));
    // ph / sh are the SpMV inputs, i.e. the vectors whose halo planes travel.  With NVLink halo push they
    // come from a process-wide pool that stays mapped into the neighbours (mapping per solve would cost
    // far more than the solve); otherwise they are ordinary buffers filled by ncclSend/Recv.
    k.push_ok = false;
    if (comm_peer_enabled() && k.halo.nspec > 0 && !getenv("RS_NO_PUSH")) {
        double *mine[2] = {nullptr, nullptr};
        k.push_ok = comm_pooled_vectors((size_t)n, mine, k.nb_lo, k.nb_hi);
        if (k.push_ok) { k.ph_v = mine[0]; k.sh_v = mine[1]; }
    }
    if (!k.push_ok) {
        RS_TRY(k.ph.alloc(n)); RS_TRY(k.sh.alloc(n));
        if (comm_active()) { RS_TRY(k.ph.zero(c.stream)); RS_TRY(k.sh.zero(c.stream)); }
        k.ph_v = k.ph.p; k.sh_v = k.sh.p;
    }
    if (precond == RS_PRECOND_JACOBI) RS_TRY(k.dinv.alloc(n));
    if (precond == RS_PRECOND_BLOCK_JACOBI) RS_TRY(k.binv.alloc((size_t)nspec * nspec * ncell));
    if (want_btd) {
        if (k.btd) { btd_destroy(k.btd); k.btd = nullptr; }
        if (is_1d) {
            const int j0 = k.halo.nspec > 0 ? k.halo.j0 : 0, j1 = k.halo.nspec > 0 ? k.halo.j1 : grid->d[0];
            RS_TRY(btd_setup(c, &k.btd, nspec, j1 - j0, grid->ncell, j0));
        } else {
            RS_TRY(btd_setup(c, &k.btd, 1, n, n, 0));
        }
    }
    if (want_mg) {
        if (k.mg) { mg_destroy(k.mg); k.mg = nullptr; }
        RS_TRY(mg_setup(c, &k.mg, *grid, k.halo.nspec > 0 ? k.halo.j0 : 0, k.halo.nspec > 0 ? k.halo.j1 : 0));
        RS_TRY(k.z.alloc(n));
        if (comm_active()) RS_TRY(k.z.zero(c.stream));
    }
    k.pool_generation = comm_pool_generation();   // after every pooled buffer of this handle has been obtained
    RS_TRY(k.partials.alloc(4 * 148 * 16));
    RS_TRY(k.scal.alloc(NSCAL));
    RS_TRY(k.ticket.alloc(4));
    RS_TRY(k.ticket.zero(c.stream));
    RS_TRY(k.scal.zero(c.stream));
    static_assert(sizeof(double) * NSCAL <= PINNED_BLOCK, "pinned block too small for the Krylov scalars");
    if (!k.h_scal) RS_CUDA(pinned_alloc((void **)&k.h_scal));
    return RS_OK;
}

// out = { max |a_ij - a_ji| , max |a_ij| } over the entries whose row and column both lie in [row0, row1);
// a missing transposed entry counts with its full size.
__global__ void __launch_bounds__(256) symmetry_kernel(int row0, int row1, const int *__restrict__ rowptr,
                                                       const int *__restrict__ col, const double *__restrict__ val,
                                                       double *partials, unsigned *ticket, double *out, int accumulate)
{
    double acc[2] = {0.0, 0.0};
    for (long long i = row0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < row1; i += (long long)gridDim.x * blockDim.x) {
        for (int p = rowptr[i]; p < rowptr[i + 1]; p++) {
            const int j = col[p];
            const double a = val[p];
            acc[1] = fmax(acc[1], fabs(a));
            if (j > i && j < row1) acc[0] = fmax(acc[0], fabs(a - csr_entry(rowptr, col, val, j, (int)i)));
            else if (j < i && j >= row0 && csr_entry(rowptr, col, val, j, (int)i) == 0.0) acc[0] = fmax(acc[0], fabs(a));
        }
    }
    grid_reduce<2>(acc, 3u, partials, ticket, out, accumulate != 0);
}

static inline double own_nnz_estimate(const Krylov &k);
// (defined with the CG kernels below)
__global__ void __launch_bounds__(256) cg_scale_setup_kernel(int row0, int row1, const double *__restrict__ dinv,
                                                             double *__restrict__ s, double *partials, unsigned *ticket,
                                                             double *out);
__global__ void __launch_bounds__(256) cg_scale_matrix_kernel(int row0, int row1, const int *__restrict__ rowptr,
                                                              const int *__restrict__ col, const double *__restrict__ s,
                                                              double sigma, int inverse, double *__restrict__ val);

// scales (or, inverse, restores) the rank's rows of the matrix in place; s must be valid on the halo planes too
static int cg_scale_matrix(Ctx &c, Krylov &k, bool inverse)
{
    for (int q = 0; q < k.own.n; q++) {
        const long long rows = k.own.e[q] - k.own.b[q];
        if (rows <= 0) continue;
        Launch L(c, K_PRECOND_SETUP, 20.0 * own_nnz_estimate(k) + 12.0 * rows);
        cg_scale_matrix_kernel<<<(unsigned)((rows * 8 + 255) / 256), 256, 0, c.stream>>>(k.own.b[q], k.own.e[q], k.rowptr, k.col, k.sc.p,
                                                                                      k.sigma, inverse ? 1 : 0, const_cast<double *>(k.val));
    }
    if (inverse) k.scaled = false;
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

// The (block-)Jacobi data of the handle's own preconditioner.  Under the multigrid cycle or the band solve nothing reads
// them, so they are only formed when one of those is not in charge -- or has just failed (krylov_solve's fallback).
static int diag_precond_setup(Ctx &c, Krylov &k)
{
    if (k.diag_valid) return RS_OK;
    if (k.precond == RS_PRECOND_JACOBI) {
        for (int q = 0; q < k.own.n; q++) {
            const int rows = k.own.e[q] - k.own.b[q];
            if (rows <= 0) continue;
            Launch L(c, K_PRECOND_SETUP, 20.0 * rows);
            jacobi_setup_kernel<<<(rows + 255) / 256, 256, 0, c.stream>>>(k.own.b[q], k.own.e[q], k.rowptr, k.col, k.val, k.dinv.p);
        }
    } else if (k.precond == RS_PRECOND_BLOCK_JACOBI) {
        const int cells = k.cell1 - k.cell0;
        if (cells > 0) {
            Launch L(c, K_PRECOND_SETUP, 20.0 * k.nspec * k.nspec * cells);
            block_setup_kernel<<<(cells + 127) / 128, 128, 0, c.stream>>>(k.nspec, k.ncell, k.cell0, k.cell1, k.rowptr, k.col, k.val, k.binv.p);
        }
    }
    RS_CUDA(cudaGetLastError());
    k.diag_valid = true;
    return RS_OK;
}

int krylov_precond_update(Ctx &c, Krylov &k)
{
    k.use_cg = false;
    k.scaled = false;
    k.norm_conv = 1.0;
    k.diag_valid = false;
    bool mg_built = false;
    // the band solve of a general structure is only decided below (btd_decide == 1): the diagonal data are formed then
    const bool ext_wanted = (k.mg != nullptr && !k.mg_failed) || (k.btd != nullptr && !k.btd_failed && k.btd_decide != 1);
    if (!ext_wanted) RS_TRY(diag_precond_setup(c, k));
    if (k.precond == RS_PRECOND_JACOBI) {
        // A Jacobian that is symmetric up to finite-difference noise (single-species diffusion-reaction: the
        // reference's own examples) is solved by preconditioned CG: one SpMV per step instead of two.
        static const bool no_cg = getenv("RS_NO_CG") != nullptr;
        if (!no_cg && k.own.n == 1 && !(k.btd && !k.btd_failed)) {
            if (k.mg != nullptr && !k.mg_failed) {
                // the hierarchy's slot-major copy of the matrix knows where every transposed entry is: build it now and
                // read the asymmetry off it (0.15 ms at 256^3 against 0.7 ms for the search through the row-wise arrays)
                RS_TRY(mg_update(c, *k.mg, k.rowptr, k.col, k.val, k.gridcols));
                mg_built = true;
                RS_TRY(mg_asymmetry(c, *k.mg, k.partials.p, k.ticket.p, k.scal.p + S_SYM));
            } else {
                Launch L(c, K_PRECOND_SETUP, 24.0 * own_nnz_estimate(k));
                symmetry_kernel<<<grid_for(k.own.e[0] - k.own.b[0], 256), 256, 0, c.stream>>>(k.own.b[0], k.own.e[0], k.rowptr, k.col, k.val,
                                                                                            k.partials.p, k.ticket.p, k.scal.p + S_SYM, 0);
            }
            RS_TRY(comm_allreduce_max(c, k.scal.p + S_SYM, 2));
            RS_CUDA(cudaMemcpyAsync(k.h_scal + S_SYM, k.scal.p + S_SYM, 2 * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
            RS_CUDA(cudaStreamSynchronize(c.stream));
            const double da = k.h_scal[S_SYM], amax = k.h_scal[S_SYM + 1];
            k.asym = amax > 0.0 ? da / amax : 1.0;
            k.use_cg = (k.asym == k.asym) && k.asym <= 1.0e-6;
            // fold the Jacobi preconditioner into the system (see the CG kernels): needs one sign on the whole diagonal
            // Under the multigrid cycle there is nothing to fold: z = M^-1 r comes from the cycle, whose smoother carries
            // the diagonal itself, and CG runs on the matrix as assembled (a negative definite J with the equally
            // negative definite cycle is CG on -J: alpha stays positive) -- no scaling pass over the matrix, no
            // unscaling of x.
            static const bool no_scale = getenv("RS_NO_CG_SCALING") != nullptr;
            const bool mg_wanted = k.mg != nullptr && !k.mg_failed;
            if (k.use_cg && !no_scale && !mg_wanted) {
                if (!k.sc.p) RS_TRY(k.sc.alloc(k.n));
                {
                    Launch L(c, K_PRECOND_SETUP, 16.0 * (double)(k.own.e[0] - k.own.b[0]));
                    cg_scale_setup_kernel<<<grid_for(k.own.e[0] - k.own.b[0], 256), 256, 0, c.stream>>>(k.own.b[0], k.own.e[0], k.dinv.p, k.sc.p,
                                                                                                      k.partials.p, k.ticket.p, k.scal.p + S_SYM);
                }
                RS_TRY(comm_allreduce_sum(c, k.scal.p + S_SYM, 3));
                RS_CUDA(cudaMemcpyAsync(k.h_scal + S_SYM, k.scal.p + S_SYM, 3 * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
                RS_CUDA(cudaStreamSynchronize(c.stream));
                const double npos = k.h_scal[S_SYM], nneg = k.h_scal[S_SYM + 1], nbad = k.h_scal[S_SYM + 2];
                if (nbad == 0.0 && (npos == 0.0 || nneg == 0.0)) {
                    k.sigma = nneg > 0.0 ? -1.0 : 1.0;
                    RS_TRY(comm_halo_exchange(c, k.sc.p, k.halo));   // columns on the neighbours' planes
                    k.scaled = true;
                    RS_TRY(cg_scale_matrix(c, k, false));
                } else {
                    k.use_cg = false;   // a diagonal of mixed sign: not a definite matrix
                }
            }
        }
    }
    // the hierarchy is rebuilt from the values the solve will use
    k.mg_on = k.mg != nullptr && !k.mg_failed;
    if (k.mg_on && !mg_built) RS_TRY(mg_update(c, *k.mg, k.rowptr, k.col, k.val, k.gridcols));
    if (k.btd && !k.btd_failed && k.btd_decide == 1) {   // general structure: is the matrix banded enough?
        double share = 0.0;
        RS_TRY(btd_band_share(c, k.n, k.rowptr, k.col, k.val, k.partials.p, k.ticket.p, k.scal.p + S_SYM, &share));
        k.btd_decide = 2;
        if (!(share >= 0.5)) k.btd_failed = true;
    }
    k.btd_on = k.btd != nullptr && !k.btd_failed;
    if (k.btd_on) RS_TRY(btd_update(c, *k.btd, k.rowptr, k.col, k.val));
    if (!k.mg_on && !k.btd_on) RS_TRY(diag_precond_setup(c, k));
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

static inline double own_nnz_estimate(const Krylov &k) { return (double)k.nnz * (double)k.own.total() / (double)k.n; }

// descriptor for pushing the first / last owned plane of vector `vi` (0: ph, 1: sh) to the slab neighbours
static HaloPush make_push(const Krylov &k, int vi, const HaloSignal &sig, bool last_launch, bool enabled = true)
{
    HaloPush hp;
    if (!k.push_ok || !enabled) return hp;
    hp.lo = (double *)k.nb_lo[vi];
    hp.hi = (double *)k.nb_hi[vi];
    hp.plane = k.halo.plane;
    hp.signal = last_launch ? 1 : 0;
    hp.sig = sig;
    return hp;
}

// One BiCGStab step; the raw dot products land in slot(cur).
static int bicgstab_step(Ctx &c, Krylov &k, double *x, int cur, int first)
{
    cudaStream_t st = c.stream;
    double *slot = k.scal.p + SLOT * cur;
    const double rows = (double)k.own.total();
    // with the multigrid cycle the vector kernels apply no preconditioner themselves (ph = M^-1 p follows them), and
    // the halo planes of ph / sh can only travel once the cycle has produced them: no push from the update kernels
    const bool ext = k.mg_on || k.btd_on;
    const int pre = ext ? RS_PRECOND_NONE : k.precond;
    const bool push = k.push_ok && !ext;
    const double pre_bytes = (pre == RS_PRECOND_BLOCK_JACOBI) ? 8.0 * k.nspec * rows : (pre == RS_PRECOND_JACOBI ? 8.0 * rows : 0.0);
    // halo handshakes of this step (sequence numbers advance identically on every rank)
    const HaloSignal sig_p = push ? comm_next_halo_signal() : HaloSignal();
    const HaloSignal sig_s = push ? comm_next_halo_signal() : HaloSignal();
    {
        Launch L(c, K_VEC, (first ? 24.0 : 40.0) * rows + pre_bytes);
        if (pre == RS_PRECOND_BLOCK_JACOBI) {
            update_p_kernel<2><<<grid_for(k.cell1 - k.cell0, 256), 256, 0, st>>>(k.cell0, k.cell1, k.nspec, k.ncell, first, k.scal.p, cur, k.r.p, k.v.p, k.p.p, k.ph_v, nullptr, k.binv.p,
                                                                                 make_push(k, 0, sig_p, true, push), k.ticket.p + 1);
        } else {
            for (int q = 0; q < k.own.n; q++) {
                const int b = k.own.b[q], e = k.own.e[q];
                if (pre == RS_PRECOND_JACOBI)
                    update_p_kernel<1><<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, 1, 0, first, k.scal.p, cur, k.r.p, k.v.p, k.p.p, k.ph_v, k.dinv.p, nullptr,
                                                                                   make_push(k, 0, sig_p, q == k.own.n - 1, push), k.ticket.p + 1);
                else
                    update_p_kernel<0><<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, 1, 0, first, k.scal.p, cur, k.r.p, k.v.p, k.p.p, k.ph_v, nullptr, nullptr,
                                                                                   make_push(k, 0, sig_p, q == k.own.n - 1, push), k.ticket.p + 1);
            }
        }
    }
    if (k.mg_on) RS_TRY(mg_apply(c, *k.mg, k.p.p, k.ph_v, nullptr, nullptr, nullptr, PeerRed()));
    if (k.btd_on) RS_TRY(btd_apply(c, *k.btd, k.p.p, k.ph_v));
    if (!push) RS_TRY(comm_halo_exchange(c, k.ph_v, k.halo));
    {
        Launch L(c, k.gridcols ? K_SPMV_GRID : K_SPMV, spmv_bytes(k, rows, own_nnz_estimate(k)) + 8.0 * rows);
        spmv_run<1>(k, st, k.ph_v, k.v.p, k.r0.p, slot + S_R0V, comm_next_peer_red(), sig_p);
    }
    if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, slot + S_R0V, 1));
    {
        Launch L(c, K_VEC, 32.0 * rows + pre_bytes);
        if (pre == RS_PRECOND_BLOCK_JACOBI) {
            update_s_kernel<2><<<grid_for(k.cell1 - k.cell0, 256), 256, 0, st>>>(k.cell0, k.cell1, k.nspec, k.ncell, k.scal.p, cur, k.r.p, k.v.p, k.s.p, k.sh_v, nullptr, k.binv.p,
                                                                                 make_push(k, 1, sig_s, true, push), k.ticket.p + 1);
        } else {
            for (int q = 0; q < k.own.n; q++) {
                const int b = k.own.b[q], e = k.own.e[q];
                if (pre == RS_PRECOND_JACOBI)
                    update_s_kernel<1><<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, 1, 0, k.scal.p, cur, k.r.p, k.v.p, k.s.p, k.sh_v, k.dinv.p, nullptr,
                                                                                   make_push(k, 1, sig_s, q == k.own.n - 1, push), k.ticket.p + 1);
                else
                    update_s_kernel<0><<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, 1, 0, k.scal.p, cur, k.r.p, k.v.p, k.s.p, k.sh_v, nullptr, nullptr,
                                                                                   make_push(k, 1, sig_s, q == k.own.n - 1, push), k.ticket.p + 1);
            }
        }
    }
    if (k.mg_on) RS_TRY(mg_apply(c, *k.mg, k.s.p, k.sh_v, nullptr, nullptr, nullptr, PeerRed()));
    if (k.btd_on) RS_TRY(btd_apply(c, *k.btd, k.s.p, k.sh_v));
    if (!push) RS_TRY(comm_halo_exchange(c, k.sh_v, k.halo));
    {
        Launch L(c, k.gridcols ? K_SPMV_GRID : K_SPMV, spmv_bytes(k, rows, own_nnz_estimate(k)) + 8.0 * rows);
        spmv_run<2>(k, st, k.sh_v, k.t.p, k.s.p, slot + S_TS, comm_next_peer_red(), sig_s);
    }
    if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, slot + S_TS, 2));
    {
        Launch L(c, K_VEC, 64.0 * rows);
        const PeerRed peer = comm_next_peer_red();
        for (int q = 0; q < k.own.n; q++) {
            const int b = k.own.b[q], e = k.own.e[q];
            update_xr_kernel<<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, k.scal.p, cur, k.ph_v, k.sh_v, k.s.p, k.t.p, k.r0.p, x, k.r.p, k.partials.p, k.ticket.p, q > 0,
                                                                         q == k.own.n - 1 ? peer : PeerRed());
        }
    }
    if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, slot + S_RR, 2));
    return RS_OK;
}

// r = r0 = b - A x explicitly; returns ||r||_2.  Restarts the recurrence (new shadow residual).
static int true_residual_restart(Ctx &c, Krylov &k, const double *b, double *x, double *rnorm)
{
    cudaStream_t st = c.stream;
    const double rows = (double)k.own.total();
    RS_TRY(comm_halo_exchange(c, x, k.halo));
    {
        Launch L(c, k.gridcols ? K_SPMV_GRID : K_SPMV, spmv_bytes(k, rows, own_nnz_estimate(k)));
        spmv_run<0>(k, st, x, k.v.p, nullptr, nullptr, PeerRed(), HaloSignal());
    }
    {
        Launch L(c, K_VEC, 32.0 * rows);
        const PeerRed peer = comm_next_peer_red();
        for (int q = 0; q < k.own.n; q++)
            krylov_restart_kernel<<<grid_for(k.own.e[q] - k.own.b[q], 256), 256, 0, st>>>(k.own.b[q], k.own.e[q], b, k.v.p, k.r.p, k.r0.p, k.partials.p, k.ticket.p, k.scal.p, q > 0,
                                                                                      q == k.own.n - 1 ? peer : PeerRed());
    }
    if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, k.scal.p + S_RTRUE, 1));
    krylov_seed_kernel<<<1, 1, 0, st>>>(k.scal.p, S_RTRUE);
    c.launches++;
    RS_CUDA(cudaMemcpyAsync(k.h_scal + S_RTRUE, k.scal.p + S_RTRUE, sizeof(double), cudaMemcpyDeviceToHost, st));
    RS_CUDA(cudaStreamSynchronize(st));
    *rnorm = std::sqrt(k.h_scal[S_RTRUE]);
    return RS_OK;
}

// ---- preconditioned CG (Jacobi) for Jacobians that are symmetric to FD noise -----------------------
// Two fused vector kernels + one SpMV per step (BiCGStab: three + two).  p lives in ph_v (the SpMV
// input whose halo planes travel), q = A p in v.  z = dinv r is never stored.

//
// The Jacobi preconditioner is folded into the system once per Newton iteration instead of being applied in every
// step: with s_i = |a_ii|^(-1/2) and sigma = sign(a_ii) (one sign for all rows, else the matrix is indefinite and
// BiCGStab takes over) CG runs on  A^ = sigma S A S  (unit diagonal, positive definite when sigma A is),
// b^ = sigma S b,  x = S x^ .  In exact arithmetic these are the iterates of Jacobi-preconditioned CG on A; the
// vector kernels no longer read dinv (16 N bytes per step) and z is r.  The values are scaled in place -- the next
// FD assembly rewrites them anyway -- and scaled back if the solve has to fall back to BiCGStab.

// s = |diag|^(-1/2) over the rows [row0,row1); out = { #rows with a_ii > 0, #rows with a_ii < 0, #rows with a_ii == 0 or NaN }
__global__ void __launch_bounds__(256) cg_scale_setup_kernel(int row0, int row1, const double *__restrict__ dinv,
                                                             double *__restrict__ s, double *partials, unsigned *ticket,
                                                             double *out)
{
    double acc[3] = {0.0, 0.0, 0.0};
    for (long long i = row0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < row1; i += (long long)gridDim.x * blockDim.x) {
        const double d = dinv[i];                 // 1 / a_ii  (jacobi_setup_kernel; 1 where a_ii is 0 or missing)
        const double a = fabs(d);
        s[i] = sqrt(a);
        if (d > 0.0 && isfinite(d)) acc[0] += 1.0; else if (d < 0.0 && isfinite(d)) acc[1] += 1.0; else acc[2] += 1.0;
    }
    grid_reduce<3>(acc, 0u, partials, ticket, out, false);
}

// val[p] *= sigma * s_i * s_col[p]   (inverse: val[p] = sigma * val[p] / (s_i * s_col[p]))
__global__ void __launch_bounds__(256) cg_scale_matrix_kernel(int row0, int row1, const int *__restrict__ rowptr,
                                                              const int *__restrict__ col, const double *__restrict__ s,
                                                              double sigma, int inverse, double *__restrict__ val)
{
    // 8 lanes per row: a row's entries are contiguous
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long i = row0 + (t >> 3);
    if (i >= row1) return;
    const double si = s[i];
    for (int p = rowptr[i] + (int)(t & 7); p < rowptr[i + 1]; p += 8) {
        const double f = si * s[col[p]];
        val[p] = inverse ? sigma * val[p] / f : sigma * f * val[p];
    }
}

// x = s * x^   (end of a scaled solve)
__global__ void __launch_bounds__(256) cg_unscale_x_kernel(int b, int e, const double *__restrict__ s, double *__restrict__ x)
{
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) x[i] *= s[i];
}

// x += alpha_prev p  (the solution update of the PREVIOUS step, deferred to here because this kernel reads the old
// p anyway: the step then touches p three times instead of four) ; p = z + beta p  [first: p = z, nothing pending],
// z = dinv r  (SCALED: z = r, dinv is not read).  A cycle ends with cg_flush_x_kernel.
template <bool SCALED>
__global__ void __launch_bounds__(256) cg_update_p_kernel(int b, int e, int first, const double *scal, int cur,
                                                          const double *__restrict__ r, const double *__restrict__ dinv,
                                                          double *__restrict__ p, double *__restrict__ x, HaloPush hp,
                                                          unsigned *hticket)
{
    double beta = 0.0, alpha = 0.0;
    if (!first) {
        beta = scal[SLOT * (cur ^ 1) + C_RZ] / scal[SLOT * cur + C_RZ];   // rho_k / rho_{k-1}
        if (!isfinite(beta)) beta = 0.0;
        alpha = scal[SLOT * (cur ^ 1) + C_ALPHA];                         // step length of the previous step (0 if it broke down)
    }
    const double2 *r2 = reinterpret_cast<const double2 *>(r), *d2 = reinterpret_cast<const double2 *>(dinv);
    double2 *p2 = reinterpret_cast<double2 *>(p), *x2 = reinterpret_cast<double2 *>(x);
#define CP_ONE(i) { double pn = (SCALED ? r[i] : dinv[i] * r[i]); \
                    if (!first) { const double po = p[i]; x[i] += alpha * po; pn += beta * po; } p[i] = pn; halo_store(hp, b, e, i, pn); }
#define CP_TWO(i) { const double2 rv = r2[i]; double2 pn = rv; if (!SCALED) { const double2 dv = d2[i]; pn.x = dv.x * rv.x; pn.y = dv.y * rv.y; } \
                    if (!first) { const double2 po = p2[i]; double2 xv = x2[i]; xv.x += alpha * po.x; xv.y += alpha * po.y; x2[i] = xv; \
                                  pn.x += beta * po.x; pn.y += beta * po.y; } p2[i] = pn; \
                    halo_store(hp, b, e, 2 * (i), pn.x); halo_store(hp, b, e, 2 * (i) + 1, pn.y); }
    RS_RANGE_LOOP(b, e, CP_ONE, CP_TWO);
#undef CP_ONE
#undef CP_TWO
    halo_signal_done(hp, hticket);
}

// x += alpha p of the last step taken (end of a cycle: before the true residual is formed, before x is handed back)
__global__ void __launch_bounds__(256) cg_flush_x_kernel(int b, int e, const double *scal, int cur, const double *__restrict__ p,
                                                         double *__restrict__ x)
{
    const double alpha = scal[SLOT * cur + C_ALPHA];
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x)
        x[i] += alpha * p[i];
}

// alpha = rho / <p,q> (kept in the slot for the deferred x update) ; r -= alpha q ; out = { <r,r>, <r, dinv r> }
// (SCALED: both are <r,r>)
template <bool SCALED>
__global__ void __launch_bounds__(256) cg_update_xr_kernel(int b, int e, double *scal, int cur, const double *__restrict__ q,
                                                           const double *__restrict__ dinv, double *__restrict__ r,
                                                           double *partials, unsigned *ticket, int accumulate, PeerRed pr)
{
    double *slot = scal + SLOT * cur;
    const double rho = scal[SLOT * (cur ^ 1) + C_RZ];
    double alpha = rho / slot[C_PQ];
    // alpha must be positive for a definite matrix (rho and <p,Ap> share the sign of A): anything else means the
    // matrix is indefinite or the recurrence broke down -- keep x (alpha = 0), tell the host.
    const bool bad = !isfinite(alpha) || !(alpha > 0.0);
    if (bad) alpha = 0.0;
    if (blockIdx.x == 0 && threadIdx.x == 0) { slot[C_BAD] = bad ? 1.0 : 0.0; slot[C_ALPHA] = alpha; }
    double acc[2] = {0.0, 0.0};
    const double2 *q2 = reinterpret_cast<const double2 *>(q), *d2 = reinterpret_cast<const double2 *>(dinv);
    double2 *r2 = reinterpret_cast<double2 *>(r);
#define CX_ONE(i) { const double rn = r[i] - alpha * q[i]; r[i] = rn; acc[0] += rn * rn; \
                    if (!SCALED) acc[1] += rn * dinv[i] * rn; }
#define CX_TWO(i) { const double2 qv = q2[i]; double2 rn = r2[i]; \
                    rn.x -= alpha * qv.x; rn.y -= alpha * qv.y; r2[i] = rn; \
                    acc[0] += rn.x * rn.x + rn.y * rn.y; \
                    if (!SCALED) { const double2 dv = d2[i]; acc[1] += rn.x * dv.x * rn.x + rn.y * dv.y * rn.y; } }
    RS_RANGE_LOOP(b, e, CX_ONE, CX_TWO);
#undef CX_ONE
#undef CX_TWO
    if (SCALED) acc[1] = acc[0];
    grid_reduce<2>(acc, 0u, partials, ticket, slot + C_RR, accumulate != 0, &pr);
}

// start of a cycle: r = b - ax (ax = nullptr: r = b, x = 0) ; out[0] = <r,r> (into scal[dst]) , out[1] = <r, dinv r>
// Scaled system (scale != nullptr): r = sigma s b - ax ; out[0] = ||r / s||^2, the residual norm of the ORIGINAL
// system, which is what the stopping test is stated in ; out[1] = <r,r>, the rho of the scaled recurrence.
__global__ void __launch_bounds__(256) cg_start_kernel(int b, int e, const double *__restrict__ rhs, const double *__restrict__ ax,
                                                       const double *__restrict__ dinv, const double *__restrict__ scale,
                                                       double sigma, double *__restrict__ r,
                                                       double *__restrict__ x, double *partials, unsigned *ticket,
                                                       double *out2, int accumulate, PeerRed pr)
{
    double acc[2] = {0.0, 0.0};
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) {
        if (scale != nullptr) {
            const double si = scale[i];
            double ri = sigma * si * rhs[i];
            if (ax != nullptr) ri -= ax[i]; else x[i] = 0.0;
            r[i] = ri;
            const double ro = ri / si;
            acc[0] += ro * ro;
            acc[1] += ri * ri;
        } else {
            double ri = rhs[i];
            if (ax != nullptr) ri -= ax[i]; else x[i] = 0.0;
            r[i] = ri;
            acc[0] += ri * ri;
            if (dinv != nullptr) acc[1] += ri * dinv[i] * ri;   // under the multigrid cycle <r, z> comes from the cycle
        }
    }
    grid_reduce<2>(acc, 0u, partials, ticket, out2, accumulate != 0, &pr);
}

// <r,r> -> where the caller reads it, <r,z> -> slot(1).rz (the "rho of the previous step" of step 0)
__global__ void cg_seed_kernel(double *scal, int dst_rr)
{
    scal[dst_rr] = scal[SLOT * 0 + 6];
    scal[SLOT * 1 + C_RZ] = scal[SLOT * 0 + 7];
}

static int cg_start(Ctx &c, Krylov &k, const double *b, double *x, bool restart, int dst_rr, double *rnorm)
{
    cudaStream_t st = c.stream;
    const double rows = (double)k.own.total();
    if (restart) {
        RS_TRY(comm_halo_exchange(c, x, k.halo));
        Launch L(c, k.gridcols ? K_SPMV_GRID : K_SPMV, spmv_bytes(k, rows, own_nnz_estimate(k)));
        spmv_run<0>(k, st, x, k.v.p, nullptr, nullptr, PeerRed(), HaloSignal());
    }
    {
        Launch L(c, K_VEC, 40.0 * rows);
        const PeerRed peer = comm_next_peer_red();
        for (int q = 0; q < k.own.n; q++)
            cg_start_kernel<<<grid_for(k.own.e[q] - k.own.b[q], 256), 256, 0, st>>>(k.own.b[q], k.own.e[q], b, restart ? k.v.p : nullptr,
                                                                                k.mg_on ? nullptr : k.dinv.p,
                                                                                k.scaled ? k.sc.p : nullptr, k.sigma,
                                                                                k.r.p, x, k.partials.p, k.ticket.p, k.scal.p + 6, q > 0,
                                                                                q == k.own.n - 1 ? peer : PeerRed());
    }
    if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, k.scal.p + 6, 2));
    cg_seed_kernel<<<1, 1, 0, st>>>(k.scal.p, dst_rr);
    c.launches++;
    if (k.mg_on) {   // preconditioned recurrence: z = M^-1 r, rho_0 = <r, z>
        RS_TRY(mg_apply(c, *k.mg, k.r.p, k.z.p, k.partials.p, k.ticket.p, k.scal.p + SLOT * 1 + C_RZ, comm_next_peer_red()));
        if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, k.scal.p + SLOT * 1 + C_RZ, 1));
    }
    RS_CUDA(cudaMemcpyAsync(k.h_scal + 6, k.scal.p + 6, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    RS_CUDA(cudaStreamSynchronize(st));
    *rnorm = std::sqrt(k.h_scal[6]);
    // what the recurrence's own <r,r> has to be multiplied by to read as a norm of the original system
    k.norm_conv = (k.scaled && k.h_scal[7] > 0.0) ? std::sqrt(k.h_scal[6] / k.h_scal[7]) : 1.0;
    return RS_OK;
}

static int cg_step(Ctx &c, Krylov &k, double *x, int cur, int first)
{
    cudaStream_t st = c.stream;
    double *slot = k.scal.p + SLOT * cur;
    const double rows = (double)k.own.total();
    const HaloSignal sig = k.push_ok ? comm_next_halo_signal() : HaloSignal();
    // the kernel variants without dinv: the Jacobi-scaled system (z = r), or z handed in by the multigrid cycle
    const bool nodinv = k.scaled || k.mg_on;
    {
        // first: r -> p ; else r, p, x -> p, x  (+ dinv when the kernel applies the Jacobi preconditioner itself)
        Launch L(c, K_VEC, ((first ? 16.0 : 40.0) + (nodinv ? 0.0 : 8.0)) * rows);
        for (int q = 0; q < k.own.n; q++) {
            const int b = k.own.b[q], e = k.own.e[q];
            if (nodinv)
                cg_update_p_kernel<true><<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, first, k.scal.p, cur, k.mg_on ? k.z.p : k.r.p, nullptr, k.ph_v, x,
                                                                                     make_push(k, 0, sig, q == k.own.n - 1), k.ticket.p + 1);
            else
                cg_update_p_kernel<false><<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, first, k.scal.p, cur, k.r.p, k.dinv.p, k.ph_v, x,
                                                                                      make_push(k, 0, sig, q == k.own.n - 1), k.ticket.p + 1);
        }
    }
    if (!k.push_ok) RS_TRY(comm_halo_exchange(c, k.ph_v, k.halo));
    {
        // <p, Ap>: the dot's second operand is the SpMV input itself -- no bytes of its own
        Launch L(c, k.gridcols ? K_SPMV_GRID : K_SPMV, spmv_bytes(k, rows, own_nnz_estimate(k)));
        spmv_run<1>(k, st, k.ph_v, k.v.p, k.ph_v, slot + C_PQ, comm_next_peer_red(), sig);
    }
    if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, slot + C_PQ, 1));
    {
        Launch L(c, K_VEC, (nodinv ? 24.0 : 32.0) * rows);   // q, r -> r (+ dinv)
        const PeerRed peer = comm_next_peer_red();
        for (int q = 0; q < k.own.n; q++) {
            const int b = k.own.b[q], e = k.own.e[q];
            if (nodinv)
                cg_update_xr_kernel<true><<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, k.scal.p, cur, k.v.p, nullptr, k.r.p,
                                                                                      k.partials.p, k.ticket.p, q > 0,
                                                                                      q == k.own.n - 1 ? peer : PeerRed());
            else
                cg_update_xr_kernel<false><<<grid_for((e - b) / 2, 256), 256, 0, st>>>(b, e, k.scal.p, cur, k.v.p, k.dinv.p, k.r.p,
                                                                                       k.partials.p, k.ticket.p, q > 0,
                                                                                       q == k.own.n - 1 ? peer : PeerRed());
        }
    }
    if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, slot + C_RR, 2));
    if (k.mg_on) {   // z = M^-1 r and the rho of the next step, <r, z>, over the <r, r> the kernel above left there
        RS_TRY(mg_apply(c, *k.mg, k.r.p, k.z.p, k.partials.p, k.ticket.p, slot + C_RZ, comm_next_peer_red()));
        if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, slot + C_RZ, 1));
    }
    return RS_OK;
}

// ---- CUDA-graph replay of Krylov steps ----------------------------------------------------------
// Small grids and sharded solves are bound by launch latency, not bandwidth: after the first two steps
// of a cycle an even number of steps (one check period; slot parities 0,1,0,1...) is captured once and
// replayed with a single host call.  Not while profiling (per-launch events cannot live in a graph) and
// not with the fused NVLink collectives (their sequence numbers are kernel arguments).
static int cg_step(Ctx &c, Krylov &k, double *x, int cur, int first);
static int bicgstab_step(Ctx &c, Krylov &k, double *x, int cur, int first);

static int graph_steps_per_launch(const Ctx &c, const Krylov &k)
{
    static const bool no_graph = getenv("RS_NO_GRAPH") != nullptr;
    if (no_graph || c.profile) return 0;
    // with the multigrid cycle a step is ~40 launches and the solve ~10 steps: on large grids launch latency is
    // nothing and a graph would only add a possibly wasted step; small grids replay two steps per host call
    // (a sharded solve is launch bound at any size: ~80 launches per step against a slab's worth of work)
    if (k.mg_on || k.btd_on) return (k.own.total() > 500000 && !comm_active() && !getenv("RS_MG_GRAPH")) ? 0 : 2;
    int g = k.check_every & ~1;
    return g >= 2 ? g : 2;
}

static int launch_step_graph(Ctx &c, Krylov &k, double *x, bool cg, int nsteps)
{
    const int w = cg ? 1 : 0;
    cudaStream_t st = c.stream;
    if (!k.g_exec[w] || k.g_x[w] != x || k.g_n[w] != nsteps || k.g_scaled[w] != k.scaled || k.g_mg[w] != (k.mg_on || k.btd_on)) {
        k.g_scaled[w] = k.scaled;
        k.g_mg[w] = k.mg_on || k.btd_on;
        if (k.g_exec[w]) { cudaGraphExecDestroy(k.g_exec[w]); k.g_exec[w] = nullptr; }
        cudaGraph_t g = nullptr;
        const long long l0 = c.launches;
        RS_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        int rc = RS_OK;
        for (int q = 0; q < nsteps && rc == RS_OK; q++) rc = cg ? cg_step(c, k, x, q & 1, 0) : bicgstab_step(c, k, x, q & 1, 0);
        cudaError_t ce = cudaStreamEndCapture(st, &g);
        k.g_launches[w] = (int)(c.launches - l0);
        c.launches = l0;
        if (rc != RS_OK) { if (g) cudaGraphDestroy(g); return rc; }
        RS_CUDA(ce);
        RS_CUDA(cudaGraphInstantiate(&k.g_exec[w], g, 0));
        cudaGraphDestroy(g);
        k.g_x[w] = x; k.g_n[w] = nsteps;
    }
    RS_CUDA(cudaGraphLaunch(k.g_exec[w], st));
    c.launches += k.g_launches[w];
    return RS_OK;
}

// PCG with the same cycle logic as the BiCGStab driver below: convergence is declared on the true residual only.
// *fallback is set when CG is not applicable after all (indefinite matrix / breakdown): the caller then runs BiCGStab.
static int pcg_solve(Ctx &c, Krylov &k, const double *b, double *x, double abs_floor, int *iters, double *relres, bool *fallback,
                     bool *converged)
{
    cudaStream_t st = c.stream;
    static const bool debug = getenv("RS_DEBUG_KRYLOV") != nullptr;
    *fallback = false;
    *converged = false;
    *iters = 0;
    *relres = 0.0;
    double bnorm = 0.0;
    RS_TRY(cg_start(c, k, b, x, false, S_BNORM, &bnorm));
    if (!(bnorm == bnorm) || std::isinf(bnorm)) { set_error("sparse solver: right-hand side is not finite"); return RS_ESINGULAR; }
    if (!(bnorm > 0.0)) { *converged = true; return RS_OK; }
    double target = k.lin_rtol * bnorm;
    if (abs_floor > target) target = abs_floor;
    double rtrue = bnorm, best = bnorm;
    int since_best = 0, cycle_it = 0, total = 0, restarts = 0, fruitless = 0;
    bool done = false;
    const int gsteps = graph_steps_per_launch(c, k);
    const int check_every = (k.mg_on || k.btd_on) ? 1 : k.check_every;   // a step under the cycle / the exact band solve is worth a look
    while (!done && total < k.max_iter) {
        int cur = cycle_it & 1;
        bool looked = false;
        if (gsteps > 0 && cycle_it >= 2 && cur == 0 && total + gsteps <= k.max_iter) {
            RS_TRY(launch_step_graph(c, k, x, true, gsteps));
            total += gsteps; cycle_it += gsteps;
            cur = 1;
            looked = true;
        } else {
            RS_TRY(cg_step(c, k, x, cur, cycle_it == 0));
            total++; cycle_it++;
        }
        *iters = total;
        if (!looked && total % check_every != 0 && total != k.max_iter) continue;
        double *slot = k.scal.p + SLOT * cur;
        RS_CUDA(cudaMemcpyAsync(k.h_scal + SLOT * cur, slot, sizeof(double) * 6, cudaMemcpyDeviceToHost, st));
        RS_CUDA(cudaStreamSynchronize(st));
        const double *h = k.h_scal + SLOT * cur;
        const double rr = h[C_RR];
        if (debug) fprintf(stderr, "[pcg] it %d rr %.6e pq %.6e rz %.6e bad %g\n", total, rr, h[C_PQ], h[C_RZ], h[C_BAD]);
        const bool finite = (rr == rr) && !std::isinf(rr);
        const bool broke = !finite || h[C_BAD] != 0.0;
        if (broke) {   // not a matrix for CG: hand over to BiCGStab (which starts again from x = 0)
            if (debug) fprintf(stderr, "[pcg] breakdown / indefinite after %d its: falling back to BiCGStab\n", total);
            *fallback = true;
            return RS_OK;
        }
        const double rnorm = std::sqrt(rr) * k.norm_conv;   // scaled solve: estimate of the original system's residual norm
        if (rnorm < 0.9 * best) { best = rnorm; since_best = 0; } else since_best++;
        const bool looks_converged = rnorm <= target;
        // CG minimises the A-norm of the error, not the residual: on an ill-conditioned matrix (a bare Laplacian,
        // condition ~ N^2) the residual sits on plateaus as long as the dimension and then drops -- a 500-box 1-D
        // transect converges in 500 steps if left alone and never if every 40-step plateau restarts the recurrence
        // (scratch check in DESIGN.md 8).  So a plateau only counts as stagnation once it is as long as the problem
        // itself (finite termination); the iteration budget bounds everything else.
        const long long plateau = std::max<long long>(40, std::min<long long>(k.n, k.max_iter));
        const bool stagnated = (long long)since_best * check_every >= plateau;
        if (!(looks_converged || stagnated || total >= k.max_iter)) continue;
        {   // the solution update of the last step is still pending (it rides on the next step's direction update)
            Launch L(c, K_VEC, 24.0 * (double)k.own.total());
            for (int q = 0; q < k.own.n; q++)
                cg_flush_x_kernel<<<grid_for(k.own.e[q] - k.own.b[q], 256), 256, 0, st>>>(k.own.b[q], k.own.e[q], k.scal.p, cur, k.ph_v, x);
        }
        double rt = 0.0;
        RS_TRY(cg_start(c, k, b, x, true, S_RTRUE, &rt));
        if (debug) fprintf(stderr, "[pcg] cycle %d ends after %d its (%s): true residual %.6e (target %.3e)\n", restarts, cycle_it,
                           looks_converged ? "converged" : stagnated ? "stagnation" : "budget", rt, target);
        if (!(rt == rt) || std::isinf(rt)) { set_error("sparse solver: Krylov breakdown (solution is not finite)"); return RS_ESINGULAR; }
        if (rt <= target) { rtrue = rt; done = true; break; }
        fruitless = (rt > 0.5 * rtrue) ? fruitless + 1 : 0;
        rtrue = rt;
        if (fruitless >= 3) { *fallback = true; return RS_OK; }
        restarts++;
        cycle_it = 0; best = rt; since_best = 0;
    }
    if (k.scaled) {   // back to the unknowns of the original system
        Launch L(c, K_VEC, 24.0 * (double)k.own.total());
        for (int q = 0; q < k.own.n; q++)
            cg_unscale_x_kernel<<<grid_for(k.own.e[q] - k.own.b[q], 256), 256, 0, st>>>(k.own.b[q], k.own.e[q], k.sc.p, x);
    }
    RS_CUDA(cudaGetLastError());
    *relres = rtrue / bnorm;
    *converged = done;
    return RS_OK;
}

// Solves A x = b from x = 0.  BiCGStab's recursively updated residual can drift away from b - A x,
// stagnate, or break down (a recurrence scalar hits zero); each of those ends the current cycle:
// the true residual is recomputed, becomes the new shadow residual, and the recurrence restarts
// from the x reached so far.  Convergence is only ever declared on the true residual.
// All ranks take the same decisions: every number they are based on has been all-reduced.
static int krylov_solve_inner(Ctx &c, Krylov &k, const double *b, double *x, double abs_floor, int *iters, double *relres, bool *converged)
{
    *converged = false;
    cudaStream_t st = c.stream;
    static const bool debug = getenv("RS_DEBUG_KRYLOV") != nullptr;
    if (k.use_cg) {
        bool fallback = false;
        RS_TRY(pcg_solve(c, k, b, x, abs_floor, iters, relres, &fallback, converged));
        if (!fallback) return RS_OK;
        *converged = false;
        k.use_cg = false;   // not a CG matrix after all: BiCGStab below, from x = 0
        if (k.scaled) {
            RS_TRY(cg_scale_matrix(c, k, true));   // ... on the matrix as assembled
            if (k.mg_on) RS_TRY(mg_update(c, *k.mg, k.rowptr, k.col, k.val, k.gridcols));   // and so is the hierarchy
        }
    }
    {
        Launch L(c, K_VEC, 32.0 * (double)k.own.total());
        const PeerRed peer = comm_next_peer_red();
        for (int q = 0; q < k.own.n; q++)
            krylov_init_kernel<<<grid_for(k.own.e[q] - k.own.b[q], 256), 256, 0, st>>>(k.own.b[q], k.own.e[q], b, k.r.p, k.r0.p, x, k.partials.p, k.ticket.p, k.scal.p, q > 0,
                                                                                   q == k.own.n - 1 ? peer : PeerRed());
    }
    if (!comm_peer_enabled()) RS_TRY(comm_allreduce_sum(c, k.scal.p + S_BNORM, 1));
    krylov_seed_kernel<<<1, 1, 0, st>>>(k.scal.p, S_BNORM);
    c.launches++;
    RS_CUDA(cudaMemcpyAsync(k.h_scal + S_BNORM, k.scal.p + S_BNORM, sizeof(double), cudaMemcpyDeviceToHost, st));
    RS_CUDA(cudaStreamSynchronize(st));
    const double bnorm = std::sqrt(k.h_scal[S_BNORM]);
    *iters = 0;
    *relres = 0.0;
    if (!(bnorm == bnorm) || std::isinf(bnorm)) { set_error("sparse solver: right-hand side is not finite"); return RS_ESINGULAR; }
    if (!(bnorm > 0.0)) { *converged = true; return RS_OK; }  // b == 0 -> x = 0
    double target = k.lin_rtol * bnorm;
    if (abs_floor > target) target = abs_floor;
    double rtrue = bnorm;     // true residual at the start of the current cycle
    double best = bnorm;      // smallest recursive residual of the current cycle
    int since_best = 0, cycle_it = 0, total = 0, restarts = 0, fruitless = 0;
    bool done = false;
    const int gsteps = graph_steps_per_launch(c, k);
    const int check_every = (k.mg_on || k.btd_on) ? 1 : k.check_every;   // a step under the cycle / the exact band solve is worth a look
    while (!done && total < k.max_iter) {
        int cur = cycle_it & 1;
        bool looked = false;
        if (gsteps > 0 && cycle_it >= 2 && cur == 0 && total + gsteps <= k.max_iter) {
            RS_TRY(launch_step_graph(c, k, x, false, gsteps));
            total += gsteps; cycle_it += gsteps;
            cur = 1;
            looked = true;   // always look after a graph: it spans a whole check period
        } else {
            RS_TRY(bicgstab_step(c, k, x, cur, cycle_it == 0));
            total++; cycle_it++;
        }
        *iters = total;
        if (!looked && total % check_every != 0 && total != k.max_iter) continue;
        double *slot = k.scal.p + SLOT * cur;
        RS_CUDA(cudaMemcpyAsync(k.h_scal + SLOT * cur, slot, sizeof(double) * SLOT, cudaMemcpyDeviceToHost, st));
        RS_CUDA(cudaStreamSynchronize(st));
        const double *h = k.h_scal + SLOT * cur;
        const double rr = h[S_RR];
        if (debug)
            fprintf(stderr, "[krylov] it %d rr %.6e r0v %.6e ts %.6e tt %.6e rhonext %.6e bad %g\n", total, rr, h[S_R0V],
                    h[S_TS], h[S_TT], h[S_RHONEXT], h[S_BAD]);
        const bool finite = (rr == rr) && !std::isinf(rr);
        const bool broke = !finite || h[S_BAD] != 0.0 || h[S_RHONEXT] == 0.0;
        const double rnorm = finite ? std::sqrt(rr) : best;
        if (finite && rnorm < 0.9 * best) { best = rnorm; since_best = 0; } else since_best++;
        const bool looks_converged = finite && !broke && rnorm <= target;
        const bool stagnated = since_best * check_every >= 40;   // ~40 steps without a 10 % gain
        if (!(looks_converged || broke || stagnated || total >= k.max_iter)) continue;
        // end of cycle: what did it really achieve?
        double rt = 0.0;
        RS_TRY(true_residual_restart(c, k, b, x, &rt));
        if (debug) fprintf(stderr, "[krylov] cycle %d ends after %d its (%s): true residual %.6e (target %.3e)\n", restarts,
                           cycle_it, looks_converged ? "converged" : broke ? "breakdown" : stagnated ? "stagnation" : "budget",
                           rt, target);
        if (!(rt == rt) || std::isinf(rt)) { set_error("sparse solver: Krylov breakdown (solution is not finite)"); return RS_ESINGULAR; }
        if (rt <= target) { rtrue = rt; done = true; break; }
        fruitless = (rt > 0.5 * rtrue) ? fruitless + 1 : 0;
        rtrue = rt;
        if (fruitless >= 6) break;   // six cycles in a row without halving the residual: give up, report what we have
        restarts++;
        cycle_it = 0; best = rt; since_best = 0;
    }
    RS_CUDA(cudaGetLastError());
    *relres = rtrue / bnorm;
    *converged = done;
    if (!done && debug) fprintf(stderr, "[krylov] not converged: relres %.3e after %d its\n", *relres, total);
    return RS_OK;
}

// The solve with the preconditioner of the handle.  Should a solve under the multigrid cycle fail -- no convergence
// within the budget, breakdown -- the handle drops the cycle for good and the system is solved again with the
// (block-)Jacobi preconditioner: never worse than without it.
int krylov_solve(Ctx &c, Krylov &k, const double *b, double *x, double abs_floor, int *iters, double *relres, bool *converged)
{
    static const bool debug = getenv("RS_DEBUG_KRYLOV") != nullptr;
    bool conv_local = false;
    if (!converged) converged = &conv_local;
    if (comm_active() && k.pool_generation != comm_pool_generation()) {
        set_error("the peer-mapped vector pool was re-created for a larger solve after this solver handle was set up; "
                  "create the handle again (one sharded solver at a time per process, or the largest one first)");
        return RS_ECOMM;
    }
    int rc = krylov_solve_inner(c, k, b, x, abs_floor, iters, relres, converged);
    if ((k.mg_on || k.btd_on) && (rc == RS_ESINGULAR || (rc == RS_OK && !*converged && !(*relres <= 0.1)))) {
        if (debug) fprintf(stderr, "[krylov] solve under the %s failed (rc %d, relres %.3e after %d its): back to Jacobi\n",
                           k.mg_on ? "multigrid cycle" : "block-tridiagonal solve", rc, *relres, *iters);
        if (k.mg_on) k.mg_failed = true;
        if (k.btd_on) k.btd_failed = true;
        k.mg_on = false;
        k.btd_on = false;
        RS_TRY(diag_precond_setup(c, k));   // skipped while the cycle / band solve was in charge
        const int it0 = *iters;
        rc = krylov_solve_inner(c, k, b, x, abs_floor, iters, relres, converged);
        *iters += it0;
    }
    return rc;
}

}  // namespace rs
