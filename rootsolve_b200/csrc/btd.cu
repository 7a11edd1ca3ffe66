// btd.cu -- exact solve of the block-tridiagonal part of the Jacobian by cyclic reduction: the preconditioner of
// the Newton linear solve on 1-D grids (steady.1D) and on general structures whose weight sits next to the diagonal.
//
// On a 1-D grid with nspec species (sparse1d, src/sparsecoeff.f:5-83) the Jacobian, seen box by box, IS block
// tridiagonal with nspec x nspec blocks -- the cell block couples the species of a box, the off-diagonal blocks the
// same species of the neighbouring boxes -- so the "preconditioner" is the reference's exact factorisation in another
// order (nsfc/nnfc, src/sparse.f:1937-2447) and the Krylov solve around it takes one or two steps: what remains for
// it are the two wrap entries of a cyclic grid, the couplings across a slab face when the solve is sharded, and
// round-off.  For a general structure (numerically detected, user supplied) the scalar tridiagonal part is used when it
// carries at least half of the off-diagonal weight: a 1-D transport model with a few non-local entries
// (man/stodes.Rd:225-252) is then solved in a handful of steps where Jacobi needs thousands.
//
// Cyclic reduction, in place: at level k (stride s = 2^k) the rows that are odd multiples of s are eliminated -- their
// inverse diagonal block is stored and they are frozen -- and the rows that are multiples of 2s absorb them:
//     alpha = A_i inv(B_{i-s}),  gamma = C_i inv(B_{i+s})
//     B_i -= alpha C_{i-s} + gamma A_{i+s},   A_i = -alpha A_{i-s},   C_i = -gamma C_{i+s}
// log2(nb) levels, every level fully parallel over its rows; alpha / gamma are kept for the right-hand sides.
// A solve is the same sweep on the vector (d_i -= alpha d_{i-s} + gamma d_{i+s}), the last row, and the levels backwards
// (x_r = inv(B_r)(d_r - A_r x_{r-s} - C_r x_{r+s})).  No pivoting across rows: stable for the diagonally dominant
// transport operators this path is made for; a solve that fails under it falls back to (block-)Jacobi (krylov.cu).
//
// Blocks are stored element-major ([e][row]) so that the threads of a level-0 launch read consecutive addresses.
#include <algorithm>
#include <cmath>
#include <memory>

#include "solver.cuh"

namespace rs {

constexpr int BTD_MAXS = 8;
constexpr int BTD_MAXLEV = 32;

struct Btd {
    int ns = 1, nb = 0;            // block size, number of block rows (of this rank)
    int stride = 0, off = 0;       // element (s, j) of a solver vector sits at s * stride + off + j
    int nlev = 0;
    long long lev_off[BTD_MAXLEV + 1] = {0};
    DevBuf<double> A, B, C, Binv;  // [ns*ns][nb]
    DevBuf<double> al, ga;         // alpha / gamma of every (kept row, level): [ns*ns][lev_off[nlev]]
    DevBuf<double> w;              // work vector [ns][nb]
};

__device__ __forceinline__ double btd_entry(const int *__restrict__ rowptr, const int *__restrict__ col,
                                            const double *__restrict__ val, long long i, long long j)
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

// the three blocks of every block row from the row-wise arrays (global indices)
__global__ void __launch_bounds__(128) btd_extract_kernel(int ns, int nb, int stride, int off, const int *__restrict__ rowptr,
                                                          const int *__restrict__ col, const double *__restrict__ val,
                                                          double *__restrict__ A, double *__restrict__ B, double *__restrict__ C)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nb) return;
    for (int s = 0; s < ns; s++) {
        const long long gi = (long long)s * stride + off + j;
        for (int s2 = 0; s2 < ns; s2++) {
            const long long gj = (long long)s2 * stride + off + j;
            const int e = s * ns + s2;
            B[(long long)e * nb + j] = btd_entry(rowptr, col, val, gi, gj);
            A[(long long)e * nb + j] = j > 0 ? btd_entry(rowptr, col, val, gi, gj - 1) : 0.0;
            C[(long long)e * nb + j] = j < nb - 1 ? btd_entry(rowptr, col, val, gi, gj + 1) : 0.0;
        }
    }
}

template <int NS>
__device__ __forceinline__ void blk_load(const double *__restrict__ M, long long nb, long long j, double (&X)[NS][NS])
{
#pragma unroll
    for (int a = 0; a < NS; a++)
#pragma unroll
        for (int b = 0; b < NS; b++) X[a][b] = M[(long long)(a * NS + b) * nb + j];
}
template <int NS>
__device__ __forceinline__ void blk_store(double *__restrict__ M, long long nb, long long j, const double (&X)[NS][NS])
{
#pragma unroll
    for (int a = 0; a < NS; a++)
#pragma unroll
        for (int b = 0; b < NS; b++) M[(long long)(a * NS + b) * nb + j] = X[a][b];
}
// Z = X Y
template <int NS>
__device__ __forceinline__ void blk_mul(const double (&X)[NS][NS], const double (&Y)[NS][NS], double (&Z)[NS][NS])
{
#pragma unroll
    for (int a = 0; a < NS; a++)
#pragma unroll
        for (int b = 0; b < NS; b++) {
            double t = 0.0;
#pragma unroll
            for (int q = 0; q < NS; q++) t += X[a][q] * Y[q][b];
            Z[a][b] = t;
        }
}

// Binv = inverse of B for the rows eliminated at this level: r = first + m * step, m = 0 .. count-1
template <int NS>
__global__ void __launch_bounds__(128) btd_invert_kernel(int nb, int first, int step, int count, const double *__restrict__ B,
                                                         double *__restrict__ Binv)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= count) return;
    const long long r = first + (long long)m * step;
    double X[NS][NS], I[NS][NS];
    blk_load<NS>(B, nb, r, X);
#pragma unroll
    for (int a = 0; a < NS; a++)
#pragma unroll
        for (int b = 0; b < NS; b++) I[a][b] = a == b ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < NS; k++) {   // Gauss-Jordan with partial pivoting inside the block
        int piv = k;
        double best = fabs(X[k][k]);
#pragma unroll
        for (int a = k + 1; a < NS; a++)
            if (fabs(X[a][k]) > best) { best = fabs(X[a][k]); piv = a; }
#pragma unroll
        for (int a = k + 1; a < NS; a++)
            if (a == piv) {
#pragma unroll
                for (int b = 0; b < NS; b++) {
                    const double tx = X[k][b]; X[k][b] = X[a][b]; X[a][b] = tx;
                    const double ti = I[k][b]; I[k][b] = I[a][b]; I[a][b] = ti;
                }
            }
        const double inv = best > 0.0 ? 1.0 / X[k][k] : 1.0;
#pragma unroll
        for (int b = 0; b < NS; b++) { X[k][b] *= inv; I[k][b] *= inv; }
#pragma unroll
        for (int a = 0; a < NS; a++) {
            if (a == k) continue;
            const double f = X[a][k];
#pragma unroll
            for (int b = 0; b < NS; b++) { X[a][b] -= f * X[k][b]; I[a][b] -= f * I[k][b]; }
        }
    }
    blk_store<NS>(Binv, nb, r, I);
}

// the rows kept at this level (i = m * 2s) absorb their neighbours i - s and i + s
template <int NS>
__global__ void __launch_bounds__(128) btd_reduce_kernel(int nb, int s, int count, double *__restrict__ A, double *__restrict__ B,
                                                         double *__restrict__ C, const double *__restrict__ Binv,
                                                         double *__restrict__ al, double *__restrict__ ga, long long nag, long long lo)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= count) return;
    const long long i = (long long)m * 2 * s, il = i - s, iu = i + s;
    double Bi[NS][NS], T[NS][NS], U[NS][NS], P[NS][NS], Z[NS][NS];
    blk_load<NS>(B, nb, i, Bi);
    // lower neighbour
    if (il >= 0) {
        blk_load<NS>(A, nb, i, T);
        blk_load<NS>(Binv, nb, il, U);
        blk_mul<NS>(T, U, P);                 // alpha
        blk_load<NS>(C, nb, il, U);
        blk_mul<NS>(P, U, Z);
#pragma unroll
        for (int a = 0; a < NS; a++)
#pragma unroll
            for (int b = 0; b < NS; b++) Bi[a][b] -= Z[a][b];
        blk_load<NS>(A, nb, il, U);
        blk_mul<NS>(P, U, Z);
#pragma unroll
        for (int a = 0; a < NS; a++)
#pragma unroll
            for (int b = 0; b < NS; b++) Z[a][b] = -Z[a][b];
        blk_store<NS>(A, nb, i, Z);
    } else {
#pragma unroll
        for (int a = 0; a < NS; a++)
#pragma unroll
            for (int b = 0; b < NS; b++) P[a][b] = 0.0;
    }
    blk_store<NS>(al, nag, lo + m, P);
    // upper neighbour
    if (iu < nb) {
        blk_load<NS>(C, nb, i, T);
        blk_load<NS>(Binv, nb, iu, U);
        blk_mul<NS>(T, U, P);                 // gamma
        blk_load<NS>(A, nb, iu, U);
        blk_mul<NS>(P, U, Z);
#pragma unroll
        for (int a = 0; a < NS; a++)
#pragma unroll
            for (int b = 0; b < NS; b++) Bi[a][b] -= Z[a][b];
        blk_load<NS>(C, nb, iu, U);
        blk_mul<NS>(P, U, Z);
#pragma unroll
        for (int a = 0; a < NS; a++)
#pragma unroll
            for (int b = 0; b < NS; b++) Z[a][b] = -Z[a][b];
        blk_store<NS>(C, nb, i, Z);
    } else {
#pragma unroll
        for (int a = 0; a < NS; a++)
#pragma unroll
            for (int b = 0; b < NS; b++) P[a][b] = 0.0;
        blk_store<NS>(C, nb, i, P);
    }
    blk_store<NS>(ga, nag, lo + m, P);
    blk_store<NS>(B, nb, i, Bi);
}

// ---- solve ---------------------------------------------------------------------------------------------------
// w(s, j) = in(s, j)   (solver vector, species-major with stride/offset, to the packed work vector [s][j])
__global__ void __launch_bounds__(256) btd_gather_kernel(int ns, int nb, int stride, int off, const double *__restrict__ in,
                                                         double *__restrict__ w)
{
    const long long total = (long long)ns * nb;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const int s = (int)(q / nb), j = (int)(q - (long long)s * nb);
        w[q] = in[(long long)s * stride + off + j];
    }
}
__global__ void __launch_bounds__(256) btd_scatter_kernel(int ns, int nb, int stride, int off, const double *__restrict__ w,
                                                          double *__restrict__ out)
{
    const long long total = (long long)ns * nb;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const int s = (int)(q / nb), j = (int)(q - (long long)s * nb);
        out[(long long)s * stride + off + j] = w[q];
    }
}

// forward: d_i -= alpha d_{i-s} + gamma d_{i+s} for the kept rows i = m * 2s
template <int NS>
__global__ void __launch_bounds__(128) btd_forward_kernel(int nb, int s, int count, const double *__restrict__ al,
                                                          const double *__restrict__ ga, long long nag, long long lo, double *__restrict__ w)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= count) return;
    const long long i = (long long)m * 2 * s, il = i - s, iu = i + s;
    double d[NS];
#pragma unroll
    for (int a = 0; a < NS; a++) d[a] = w[(long long)a * nb + i];
    if (il >= 0) {
        double P[NS][NS];
        blk_load<NS>(al, nag, lo + m, P);
#pragma unroll
        for (int a = 0; a < NS; a++)
#pragma unroll
            for (int b = 0; b < NS; b++) d[a] -= P[a][b] * w[(long long)b * nb + il];
    }
    if (iu < nb) {
        double P[NS][NS];
        blk_load<NS>(ga, nag, lo + m, P);
#pragma unroll
        for (int a = 0; a < NS; a++)
#pragma unroll
            for (int b = 0; b < NS; b++) d[a] -= P[a][b] * w[(long long)b * nb + iu];
    }
#pragma unroll
    for (int a = 0; a < NS; a++) w[(long long)a * nb + i] = d[a];
}

// backward: x_r = inv(B_r) (d_r - A_r x_{r-s} - C_r x_{r+s}) for the rows r = first + m * step eliminated at stride s
// (s = 0: the last remaining row, no neighbours)
template <int NS>
__global__ void __launch_bounds__(128) btd_backward_kernel(int nb, int first, int step, int s, int count, const double *__restrict__ A,
                                                           const double *__restrict__ C, const double *__restrict__ Binv,
                                                           double *__restrict__ w)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= count) return;
    const long long r = first + (long long)m * step, rl = r - s, ru = r + s;
    double d[NS], x[NS];
#pragma unroll
    for (int a = 0; a < NS; a++) d[a] = w[(long long)a * nb + r];
    if (s > 0) {
        double P[NS][NS];
        if (rl >= 0) {
            blk_load<NS>(A, nb, r, P);
#pragma unroll
            for (int a = 0; a < NS; a++)
#pragma unroll
                for (int b = 0; b < NS; b++) d[a] -= P[a][b] * w[(long long)b * nb + rl];
        }
        if (ru < nb) {
            blk_load<NS>(C, nb, r, P);
#pragma unroll
            for (int a = 0; a < NS; a++)
#pragma unroll
                for (int b = 0; b < NS; b++) d[a] -= P[a][b] * w[(long long)b * nb + ru];
        }
    }
    double I[NS][NS];
    blk_load<NS>(Binv, nb, r, I);
#pragma unroll
    for (int a = 0; a < NS; a++) {
        double t = 0.0;
#pragma unroll
        for (int b = 0; b < NS; b++) t += I[a][b] * d[b];
        x[a] = t;
    }
#pragma unroll
    for (int a = 0; a < NS; a++) w[(long long)a * nb + r] = x[a];
}

// out = { sum |a_ij| over |i - j| = 1 , sum |a_ij| over i != j }: how much of the coupling the tridiagonal part carries
__global__ void __launch_bounds__(256) btd_band_weight_kernel(int n, const int *__restrict__ rowptr, const int *__restrict__ col,
                                                              const double *__restrict__ val, double *partials, unsigned *ticket,
                                                              double *out)
{
    double acc[2] = {0.0, 0.0};
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        for (int p = rowptr[i]; p < rowptr[i + 1]; p++) {
            const long long j = col[p];
            if (j == i) continue;
            const double a = fabs(val[p]);
            acc[1] += a;
            if (j == i - 1 || j == i + 1) acc[0] += a;
        }
    grid_reduce<2>(acc, 0u, partials, ticket, out, false);
}

// ---- host side -------------------------------------------------------------------------------------------------
void btd_destroy(Btd *b) { delete b; }

int btd_setup(Ctx &c, Btd **out, int ns, int nb, int stride, int off)
{
    if (ns < 1 || ns > BTD_MAXS || nb < 1) { set_error("block-tridiagonal preconditioner: unsupported block size %d", ns); return RS_EINVAL; }
    std::unique_ptr<Btd> T(new Btd());
    T->ns = ns; T->nb = nb; T->stride = stride; T->off = off;
    int L = 0;
    while ((1LL << L) < nb) L++;
    T->nlev = L;
    long long tot = 0;
    for (int k = 0; k < L; k++) {
        T->lev_off[k] = tot;
        tot += (nb + (2LL << k) - 1) / (2LL << k);   // rows that are multiples of 2^(k+1)
    }
    T->lev_off[L] = tot;
    const size_t e = (size_t)ns * ns;
    RS_TRY(T->A.alloc(e * nb)); RS_TRY(T->B.alloc(e * nb)); RS_TRY(T->C.alloc(e * nb)); RS_TRY(T->Binv.alloc(e * nb));
    RS_TRY(T->al.alloc(e * (size_t)std::max<long long>(tot, 1))); RS_TRY(T->ga.alloc(e * (size_t)std::max<long long>(tot, 1)));
    RS_TRY(T->w.alloc((size_t)ns * nb));
    (void)c;
    *out = T.release();
    return RS_OK;
}

template <int NS>
static void btd_factor_t(Ctx &c, Btd &T)
{
    cudaStream_t st = c.stream;
    const int nb = T.nb;
    const long long nag = std::max<long long>(T.lev_off[T.nlev], 1);
    for (int k = 0; k < T.nlev; k++) {
        const int s = 1 << k;
        const int nel = (nb - s + 2 * s - 1) / (2 * s);        // odd multiples of s below nb
        const int nkeep = (int)(T.lev_off[k + 1] - T.lev_off[k]);
        if (nel > 0) {
            Launch L(c, K_PRECOND_SETUP, 16.0 * NS * NS * nel);
            btd_invert_kernel<NS><<<(nel + 127) / 128, 128, 0, st>>>(nb, s, 2 * s, nel, T.B.p, T.Binv.p);
        }
        Launch L(c, K_PRECOND_SETUP, 80.0 * NS * NS * nkeep);
        btd_reduce_kernel<NS><<<(nkeep + 127) / 128, 128, 0, st>>>(nb, s, nkeep, T.A.p, T.B.p, T.C.p, T.Binv.p, T.al.p, T.ga.p, nag, T.lev_off[k]);
    }
    Launch L(c, K_PRECOND_SETUP, 16.0 * NS * NS);
    btd_invert_kernel<NS><<<1, 128, 0, st>>>(nb, 0, 1, 1, T.B.p, T.Binv.p);   // the last remaining row
}

template <int NS>
static void btd_solve_t(Ctx &c, Btd &T)
{
    cudaStream_t st = c.stream;
    const int nb = T.nb;
    const long long nag = std::max<long long>(T.lev_off[T.nlev], 1);
    for (int k = 0; k < T.nlev; k++) {
        const int s = 1 << k;
        const int nkeep = (int)(T.lev_off[k + 1] - T.lev_off[k]);
        Launch L(c, K_PRECOND_APPLY, (16.0 * NS * NS + 32.0 * NS) * nkeep);
        btd_forward_kernel<NS><<<(nkeep + 127) / 128, 128, 0, st>>>(nb, s, nkeep, T.al.p, T.ga.p, nag, T.lev_off[k], T.w.p);
    }
    {
        Launch L(c, K_PRECOND_APPLY, 8.0 * NS * NS + 16.0 * NS);
        btd_backward_kernel<NS><<<1, 128, 0, st>>>(nb, 0, 1, 0, 1, T.A.p, T.C.p, T.Binv.p, T.w.p);
    }
    for (int k = T.nlev - 1; k >= 0; k--) {
        const int s = 1 << k;
        const int nel = (nb - s + 2 * s - 1) / (2 * s);
        if (nel <= 0) continue;
        Launch L(c, K_PRECOND_APPLY, (24.0 * NS * NS + 32.0 * NS) * nel);
        btd_backward_kernel<NS><<<(nel + 127) / 128, 128, 0, st>>>(nb, s, 2 * s, s, nel, T.A.p, T.C.p, T.Binv.p, T.w.p);
    }
}

#define BTD_DISPATCH(FN, ...)                     \
    switch (T.ns) {                               \
    case 1: FN<1>(__VA_ARGS__); break;            \
    case 2: FN<2>(__VA_ARGS__); break;            \
    case 3: FN<3>(__VA_ARGS__); break;            \
    case 4: FN<4>(__VA_ARGS__); break;            \
    case 5: FN<5>(__VA_ARGS__); break;            \
    case 6: FN<6>(__VA_ARGS__); break;            \
    case 7: FN<7>(__VA_ARGS__); break;            \
    default: FN<8>(__VA_ARGS__); break;           \
    }

// factorisation of the block-tridiagonal part of the current matrix (once per Newton iteration)
int btd_update(Ctx &c, Btd &T, const int *rowptr, const int *col, const double *val)
{
    {
        Launch L(c, K_PRECOND_SETUP, 36.0 * T.ns * T.ns * T.nb);
        btd_extract_kernel<<<(T.nb + 127) / 128, 128, 0, c.stream>>>(T.ns, T.nb, T.stride, T.off, rowptr, col, val, T.A.p, T.B.p, T.C.p);
    }
    BTD_DISPATCH(btd_factor_t, c, T);
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

// z = T^-1 r on the rank's block rows of global-index vectors
int btd_apply(Ctx &c, Btd &T, const double *r, double *z)
{
    const long long tot = (long long)T.ns * T.nb;
    {
        Launch L(c, K_PRECOND_APPLY, 16.0 * tot);
        btd_gather_kernel<<<grid_for(tot, 256), 256, 0, c.stream>>>(T.ns, T.nb, T.stride, T.off, r, T.w.p);
    }
    BTD_DISPATCH(btd_solve_t, c, T);
    {
        Launch L(c, K_PRECOND_APPLY, 16.0 * tot);
        btd_scatter_kernel<<<grid_for(tot, 256), 256, 0, c.stream>>>(T.ns, T.nb, T.stride, T.off, T.w.p, z);
    }
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

// share of the off-diagonal weight of the rows [0, n) that lies on the two neighbouring diagonals (host value; one sync)
int btd_band_share(Ctx &c, int n, const int *rowptr, const int *col, const double *val, double *partials, unsigned *ticket,
                   double *d_out2, double *share)
{
    {
        Launch L(c, K_PRECOND_SETUP, 12.0 * 3.0 * n);
        btd_band_weight_kernel<<<grid_for(n, 256), 256, 0, c.stream>>>(n, rowptr, col, val, partials, ticket, d_out2);
    }
    double h[2] = {0.0, 0.0};
    RS_CUDA(cudaMemcpyAsync(h, d_out2, 2 * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    *share = h[1] > 0.0 ? h[0] / h[1] : 0.0;
    return RS_OK;
}

}  // namespace rs
