// fdjac.cu -- the grouped finite-difference Jacobian (replaces xSparseJacob,
// src/sparse.f:461-525).
//
// The reference perturbs the columns of one group in place, calls the model,
// harvests the group's columns, restores, and moves to the next group: ngp+1
// sequential model calls.  Here the ngp+1 state vectors are laid out side by
// side (K_PERTURB), ALL of them are evaluated by one batched launch of the
// model's residual (K_RESIDUAL, grid.y = evaluation), and one assembly pass
// (K_ASSEMBLE) writes every entry of J straight into the row-wise arrays the
// Krylov solver reads:
//     J(i,j) = (f_g(j)[i] - f_0[i]) / ((y_j + delta_j) - y_j)
// with delta = max(|y_j| 1e-8, 1e-8) (Perturb, src/steady.f:219-224) and the
// ROUNDED difference as divisor (src/sparse.f:511).  beta = -f_0.
// The arithmetic is ordered as in the reference and this file is compiled with
// -fmad=false, so a[] is bit-identical to the reference's when the model's own
// arithmetic is.
#include <algorithm>
#include <cstdlib>

#include "grid_stencil.cuh"
#include "solver.cuh"

namespace rs {

__device__ __forceinline__ double perturbed(double v)
{
    double d = fabs(v) * 1.0e-8;
    if (d < 1.0e-8) d = 1.0e-8;
    return v + d;
}

// Y[g+1][i] = y[i] (+delta_i when column i belongs to group g), for i in [b, e); plane 0 is not used (f(y) is
// evaluated on y itself, fd_base)
__global__ void __launch_bounds__(256) perturb_kernel(int b, int e, int ngp, const double *__restrict__ y,
                                                      const int *__restrict__ group_of, double *__restrict__ Y,
                                                      long long ld)
{
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) {
        const double v = y[i];
        const double vp = perturbed(v);
        const int g = group_of[i];
        for (int q = 0; q < ngp; q++) Y[(long long)(q + 1) * ld + i] = (q == g) ? vp : v;
    }
}

// LANES threads per row; entry p of row i: column j = col[p].
template <int LANES>
__global__ void __launch_bounds__(256) assemble_kernel(int row0, int row1, const int *__restrict__ rowptr,
                                                       const int *__restrict__ col, const int *__restrict__ group_of,
                                                       const double *__restrict__ y, const double *__restrict__ F,
                                                       long long ld, double *__restrict__ val,
                                                       double *__restrict__ beta)
{
    const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = (int)(gt % LANES);
    const long long nrows = row1 - row0;
    for (long long r = gt / LANES; r < nrows; r += ((long long)gridDim.x * blockDim.x) / LANES) {
        const int i = row0 + (int)r;
        const double f0 = F[i];
        if (lane == 0) beta[i] = -f0;
        const int e = rowptr[i + 1];
        for (int p = rowptr[i] + lane; p < e; p += LANES) {
            const int j = col[p];
            const double yj = y[j];
            const double divdelt = perturbed(yj) - yj;
            const double fg = F[(long long)(group_of[j] + 1) * ld + i];
            val[p] = (fg + (-f0)) / divdelt;
        }
    }
}

// The same on a verified regular grid (FdWork::gridcols): one thread per row, the row's columns computed from its
// grid position (GridRow, ascending = the order of the row-wise arrays), so col[] is not read; all gathers of a row
// (y, group, then the residual planes) are issued back to back, and the values of a 256-row tile leave through shared
// memory as one contiguous, coalesced slice of val[].  Same arithmetic per entry as assemble_kernel.
template <int NSPEC>
__global__ void __launch_bounds__(256) assemble_grid_kernel(GridCols g, int row0, int row1, const int *__restrict__ rowptr,
                                                            const int *__restrict__ group_of, const double *__restrict__ y,
                                                            const double *__restrict__ F, long long ld,
                                                            double *__restrict__ val, double *__restrict__ beta)
{
    constexpr int NSLOT = GridRow<NSPEC>::NSLOT;
    __shared__ double sv[256 * NSLOT];
    __shared__ int s_base, s_end;
    const int ntiles = (row1 - row0 + 255) / 256;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int i0 = row0 + tile * 256, i1 = min(row1, i0 + 256);
        const int i = i0 + (int)threadIdx.x;
        const bool valid = i < i1;
        if (threadIdx.x == 0) { s_base = rowptr[i0]; s_end = rowptr[i1]; }
        GridRow<NSPEC> gr;
        gr.set(g, valid ? i : i0, valid);
        const double f0 = valid ? F[i] : 0.0;
        if (valid) beta[i] = -f0;
        const int pb = valid ? rowptr[i] : 0;
        double yv[NSLOT], fg[NSLOT];
        int gg[NSLOT];
#pragma unroll
        for (int q = 0; q < NSLOT; q++) {
            yv[q] = gr.has[q] ? y[gr.c[q]] : 0.0;
            gg[q] = gr.has[q] ? group_of[gr.c[q]] : 0;
        }
#pragma unroll
        for (int q = 0; q < NSLOT; q++) fg[q] = gr.has[q] ? F[(long long)(gg[q] + 1) * ld + i] : 0.0;
        __syncthreads();
        int p = pb - s_base;
#pragma unroll
        for (int q = 0; q < NSLOT; q++)
            if (gr.has[q]) {
                const double divdelt = perturbed(yv[q]) - yv[q];
                sv[p++] = (fg[q] + (-f0)) / divdelt;
            }
        __syncthreads();
        const int cnt = s_end - s_base;
        for (int e = threadIdx.x; e < cnt; e += 256) val[s_base + e] = sv[e];
        __syncthreads();
    }
}

int fd_setup(Ctx &c, FdWork &w, int n, int ngp, int nout, const double *rpar, int nrpar, const int *ipar,
             int nipar, int slab_j0, int slab_j1)
{
    // 256-byte aligned leading dimension so every evaluation starts on a fresh line
    w.ld = ((long long)n + 31) / 32 * 32;
    w.ngp_alloc = ngp;
    // Y (the ngp perturbed states side by side) is only needed by models in pointer form, the ngp residual planes of
    // F only by models without a row function: both are allocated on first use; plane 0 of F holds f(y)
    RS_TRY(w.F.alloc((size_t)w.ld));
    // ipar / out exactly as initOut builds them (src/steady_utils.c:109-124), then the slab bounds
    const int lrpar = nout + nrpar, lipar = 3 + nipar;
    w.ldout = lrpar > 0 ? lrpar : 1;
    std::vector<int> ip((size_t)lipar + 2);
    ip[0] = nout; ip[1] = lrpar; ip[2] = lipar;
    for (int q = 0; q < nipar; q++) ip[3 + q] = ipar[q];
    ip[lipar] = slab_j0; ip[lipar + 1] = slab_j1;
    RS_TRY(w.ip.alloc(ip.size()));
    RS_TRY(w.ip.upload(ip.data(), ip.size(), c.stream));
    std::vector<double> out((size_t)w.ldout * (ngp + 1), 0.0);
    for (int b = 0; b <= ngp; b++)
        for (int q = 0; q < nrpar; q++) out[(size_t)b * w.ldout + nout + q] = rpar[q];
    RS_TRY(w.yout.alloc(out.size()));
    RS_TRY(w.yout.upload(out.data(), out.size(), c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    return RS_OK;
}

// beta = -f(y) alone (the first model call of xSparseJacob, src/sparse.f:487-491).  The Newton loop runs it first:
// convergence test 1 (src/sparse.f:122-133) needs nothing else, and when the test passes the ngp perturbed
// evaluations of that iteration -- whose Jacobian the reference computes and never uses -- are not made.
__global__ void __launch_bounds__(256) negate_kernel(int b, int e, const double *__restrict__ f, double *__restrict__ beta)
{
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) beta[i] = -f[i];
}

int fd_base(Ctx &c, const rs_model_desc *m, FdWork &w, int n, double t, const double *d_y, double *d_beta, const Ranges &own)
{
    const long long work = own.total();
    {
        Launch L(c, K_RESIDUAL, 16.0 * work);
        const int e = m->launch((void *)c.stream, n, t, d_y, w.F.p, w.ld, 1, w.yout.p, w.ldout, w.ip.p, work);
        if (e) { set_error("model launch failed: %s", cudaGetErrorString((cudaError_t)e)); return RS_ENODEVICE; }
    }
    {
        Launch L(c, K_NEWTON, 16.0 * work);
        for (int q = 0; q < own.n; q++)
            negate_kernel<<<grid_for(own.e[q] - own.b[q], 256), 256, 0, c.stream>>>(own.b[q], own.e[q], w.F.p, d_beta);
    }
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

// The ngp perturbed evaluations and the assembly; F[0] = f(y) is already there (fd_base).
int fd_groups(Ctx &c, const rs_model_desc *m, FdWork &w, int n, int nnz, int ngp, double t, const double *d_y,
              const int *d_group_of, const int *d_rowptr, const int *d_col, double *d_val, double *d_beta,
              const Ranges &own, const Ranges &touch)
{
    const long long work = own.total();
    const double frac = (double)work / (double)n;
    const bool no_rows = getenv("RS_NO_FD_ROWS") != nullptr;   // tests compare the two routes
    if (m->launch_fd_rows && !no_rows) {
        // row form: every entry of J from one evaluation of the model's row function with the entry's column perturbed;
        // bit-identical to the grouped evaluation (a column of a group is the only one of its group that reaches the row)
        Launch L(c, K_FD_ROWS, (8.0 * n + 4.0 * (n + 1) + 12.0 * nnz) * frac);
        for (int q = 0; q < own.n; q++) {
            const int e = m->launch_fd_rows((void *)c.stream, own.b[q], own.e[q], t, d_y, d_rowptr, d_col, d_val, w.yout.p, w.ip.p);
            if (e) { set_error("model launch failed: %s", cudaGetErrorString((cudaError_t)e)); return RS_ENODEVICE; }
        }
        return RS_OK;
    }
    if (w.F.n < (size_t)w.ld * (w.ngp_alloc + 1)) {   // grow F to its ngp + 1 planes, keeping f(y) in plane 0
        DevBuf<double> big;
        RS_TRY(big.alloc((size_t)w.ld * (w.ngp_alloc + 1)));
        RS_CUDA(cudaMemcpyAsync(big.p, w.F.p, sizeof(double) * (size_t)w.ld, cudaMemcpyDeviceToDevice, c.stream));
        RS_CUDA(cudaStreamSynchronize(c.stream));
        std::swap(w.F.p, big.p);
        std::swap(w.F.n, big.n);
    }
    if (m->launch_groups && ngp > 0) {
        // accessor-form model: the ngp perturbed states are never written -- every evaluation reads y and the group
        // map and adds its increment on the fly (B_res = 8N + 4N + 8N per evaluation written)
        Launch L(c, K_RESIDUAL, 12.0 * work + 8.0 * work * ngp);
        const int e = m->launch_groups((void *)c.stream, n, t, d_y, d_group_of, 0, ngp, w.F.p + w.ld, w.ld, w.yout.p + w.ldout, w.ldout,
                                       w.ip.p, work);
        if (e) { set_error("model launch failed: %s", cudaGetErrorString((cudaError_t)e)); return RS_ENODEVICE; }
    } else if (ngp > 0) {
        if (!w.Y.p) RS_TRY(w.Y.alloc((size_t)w.ld * (w.ngp_alloc + 1)));
        {
            const double nt = (double)touch.total();
            Launch L(c, K_PERTURB, 12.0 * nt + 8.0 * nt * ngp);
            for (int q = 0; q < touch.n; q++)
                perturb_kernel<<<grid_for(touch.e[q] - touch.b[q], 256), 256, 0, c.stream>>>(touch.b[q], touch.e[q], ngp, d_y, d_group_of,
                                                                                           w.Y.p, w.ld);
        }
        Launch L(c, K_RESIDUAL, 16.0 * work * ngp);
        const int e = m->launch((void *)c.stream, n, t, w.Y.p + w.ld, w.F.p + w.ld, w.ld, ngp, w.yout.p + w.ldout, w.ldout, w.ip.p, work);
        if (e) { set_error("model launch failed: %s", cudaGetErrorString((cudaError_t)e)); return RS_ENODEVICE; }
    }
    {
        Launch L(c, K_ASSEMBLE, (20.0 * nnz + 20.0 * n) * frac);
        for (int q = 0; q < own.n; q++) {
            const long long rows = own.e[q] - own.b[q];
            if (rows <= 0) continue;
            if (w.gridcols) {
                const GridCols g{w.gd1, w.gd2, w.gd3, w.gd2 * w.gd3, w.gncell};
                const int blocks = (int)std::min<long long>((rows + 255) / 256, 148 * 8);
                if (w.gridcols == 1)
                    assemble_grid_kernel<1><<<blocks, 256, 0, c.stream>>>(g, own.b[q], own.e[q], d_rowptr, d_group_of, d_y, w.F.p, w.ld, d_val, d_beta);
                else
                    assemble_grid_kernel<2><<<blocks, 256, 0, c.stream>>>(g, own.b[q], own.e[q], d_rowptr, d_group_of, d_y, w.F.p, w.ld, d_val, d_beta);
            } else {
                assemble_kernel<8><<<grid_for(rows * 8, 256, 148 * 16), 256, 0, c.stream>>>(own.b[q], own.e[q], d_rowptr, d_col,
                                                                                          d_group_of, d_y, w.F.p, w.ld, d_val, d_beta);
            }
        }
    }
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

int fd_jacobian(Ctx &c, const rs_model_desc *m, FdWork &w, int n, int nnz, int ngp, double t, const double *d_y,
                const int *d_group_of, const int *d_rowptr, const int *d_col, double *d_val, double *d_beta,
                const Ranges &own, const Ranges &touch)
{
    RS_TRY(fd_base(c, m, w, n, t, d_y, d_beta, own));
    return fd_groups(c, m, w, n, nnz, ngp, t, d_y, d_group_of, d_rowptr, d_col, d_val, d_beta, own, touch);
}

int residual_only(Ctx &c, const rs_model_desc *m, FdWork &w, int n, double t, const double *d_y, double *d_f,
                  long long work)
{
    Launch L(c, K_RESIDUAL, 16.0 * work);
    const int e = m->launch((void *)c.stream, n, t, d_y, d_f, w.ld, 1, w.yout.p, w.ldout, w.ip.p, work);
    if (e) { set_error("model launch failed: %s", cudaGetErrorString((cudaError_t)e)); return RS_ENODEVICE; }
    return RS_OK;
}

// ---- numeric structure detection (Type = +-1, sparsetype "sparseint"/"sparsereturn") -------------------
// The reference perturbs one state variable at a time and records which rates of change move
// (src/sparse.f:356-396): N+1 model calls.  Here the N perturbed states are evaluated in batches of B by
// the batched residual kernel; a bit matrix (column j, row i) remembers the hits, so the model is not
// called a second time to fill jan.  Same test as the reference: |(f_j[i] - f[i]) / delta_j| > 1e-30 with
// delta_j the increment Perturb returns (not the rounded difference); rows come out ascending.

// Y[b][i] = y[i] (+ delta when i == j0 + b), b in [0, nb)
__global__ void __launch_bounds__(256) perturb_single_kernel(int n, int j0, int nb, const double *__restrict__ y,
                                                             double *__restrict__ Y, long long ld)
{
    const long long total = (long long)n * nb;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(q / n), i = (int)(q - (long long)b * n);
        const double v = y[i];
        Y[(long long)b * ld + i] = (i == j0 + b) ? perturbed(v) : v;
    }
}

__global__ void __launch_bounds__(256) detect_kernel(int n, int j0, int nb, const double *__restrict__ y,
                                                     const double *__restrict__ F0, const double *__restrict__ F,
                                                     long long ld, unsigned *__restrict__ mask, int words,
                                                     int *__restrict__ counts)
{
    const long long total = (long long)words * nb;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(q / words), w = (int)(q - (long long)b * words);
        const int j = j0 + b;
        const double yj = y[j];
        double delta = fabs(yj) * 1.0e-8;
        if (delta < 1.0e-8) delta = 1.0e-8;
        unsigned bits = 0;
        for (int t = 0; t < 32; t++) {
            const int i = w * 32 + t;
            if (i < n && fabs((F[(long long)b * ld + i] - F0[i]) / delta) > 1.0e-30) bits |= 1u << t;
        }
        mask[(size_t)j * words + w] = bits;
        if (bits) atomicAdd(&counts[j], __popc(bits));
    }
}

__global__ void detect_fill_kernel(int n, const unsigned *__restrict__ mask, int words, const int *__restrict__ ian,
                                   int *__restrict__ jan)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int pos = ian[j] - 1;
    for (int w = 0; w < words; w++) {
        unsigned bits = mask[(size_t)j * words + w];
        while (bits) {
            const int t = __ffs(bits) - 1;
            bits &= bits - 1;
            jan[pos++] = w * 32 + t + 1;
        }
    }
}

int detect_structure(Ctx &c, const rs_model_desc *m, int n, double t, const double *d_y, int nout, const double *rpar,
                     int nrpar, const int *ipar, int nipar, int nnz_cap, DevBuf<int> &ian, DevBuf<int> &jan, int *nnz)
{
    const long long ld = ((long long)n + 31) / 32 * 32;
    long long B = (1LL << 27) / ld;   // <= 1 GiB per work array
    if (B < 1) B = 1;
    if (B > n) B = n;
    FdWork w;
    RS_TRY(fd_setup(c, w, n, (int)B - 1, nout, rpar, nrpar, ipar, nipar, 0, -1));
    RS_TRY(w.Y.alloc((size_t)w.ld * B));
    RS_TRY(w.F.alloc((size_t)w.ld * B));
    const int words = (n + 31) / 32;
    DevBuf<double> F0;
    DevBuf<unsigned> mask;
    DevBuf<int> counts;
    RS_TRY(F0.alloc(n));
    RS_TRY(mask.alloc((size_t)n * words));
    RS_TRY(counts.alloc(n));
    RS_TRY(counts.zero(c.stream));
    RS_TRY(residual_only(c, m, w, n, t, d_y, F0.p, n));
    for (int j0 = 0; j0 < n; j0 += (int)B) {
        const int nb = (int)std::min<long long>(B, n - j0);
        {
            Launch L(c, K_PERTURB, 8.0 * n * (nb + 1));
            perturb_single_kernel<<<grid_for((long long)n * nb, 256), 256, 0, c.stream>>>(n, j0, nb, d_y, w.Y.p, w.ld);
        }
        {
            Launch L(c, K_RESIDUAL, 16.0 * n * nb);
            const int e = m->launch((void *)c.stream, n, t, w.Y.p, w.F.p, w.ld, nb, w.yout.p, w.ldout, w.ip.p, n);
            if (e) { set_error("model launch failed: %s", cudaGetErrorString((cudaError_t)e)); return RS_ENODEVICE; }
        }
        {
            Launch L(c, K_STRUCT, 8.0 * n * nb);
            detect_kernel<<<grid_for((long long)words * nb, 256), 256, 0, c.stream>>>(n, j0, nb, d_y, F0.p, w.F.p, w.ld, mask.p, words, counts.p);
        }
    }
    RS_TRY(ian.alloc((size_t)n + 1));
    RS_TRY(exclusive_scan(c, counts.p, ian.p, n, 1, true));
    int last = 0;
    RS_CUDA(cudaMemcpyAsync(&last, ian.p + n, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    *nnz = last - 1;
    if (*nnz > nnz_cap) {   // src/sparse.f:374-393
        set_error("error during determining sparsity: nnz too small; nnz should be at least %d", *nnz);
        return RS_ESTRUCT;
    }
    RS_TRY(jan.alloc((size_t)std::max(*nnz, 1)));
    {
        Launch L(c, K_STRUCT, 4.0 * *nnz + 4.0 * (double)n * words);
        detect_fill_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(n, mask.p, words, ian.p, jan.p);
    }
    RS_CUDA(cudaGetLastError());
    RS_CUDA(cudaStreamSynchronize(c.stream));
    return RS_OK;
}

// ---- jan-order <-> row-wise values --------------------------------------------
__device__ __forceinline__ int find_in_row(const int *__restrict__ col, int b, int e, int j)
{
    while (b < e) {
        const int mid = (b + e) >> 1;
        const int cj = col[mid];
        if (cj == j) return mid;
        if (cj < j) b = mid + 1; else e = mid;
    }
    return -1;
}

template <bool TO_CSC>
__global__ void csc_csr_kernel(int n, const int *__restrict__ ian, const int *__restrict__ jan,
                               const int *__restrict__ rowptr, const int *__restrict__ col, double *__restrict__ val,
                               double *__restrict__ a)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    for (int k = ian[j] - 1; k < ian[j + 1] - 1; k++) {
        const int i = jan[k] - 1;
        const int p = find_in_row(col, rowptr[i], rowptr[i + 1], j);
        if (p < 0) continue;
        if (TO_CSC) a[k] = val[p]; else val[p] = a[k];
    }
}

int gather_csc_values(Ctx &c, int n, int nnz, const int *d_ian, const int *d_jan, const int *d_rowptr,
                      const int *d_col, const double *d_val, double *d_a)
{
    Launch L(c, K_MISC, 24.0 * nnz);
    csc_csr_kernel<true><<<(n + 255) / 256, 256, 0, c.stream>>>(n, d_ian, d_jan, d_rowptr, d_col,
                                                                const_cast<double *>(d_val), d_a);
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

int scatter_csc_values(Ctx &c, int n, int nnz, const int *d_ian, const int *d_jan, const int *d_rowptr,
                       const int *d_col, const double *d_a, double *d_val)
{
    Launch L(c, K_MISC, 24.0 * nnz);
    csc_csr_kernel<false><<<(n + 255) / 256, 256, 0, c.stream>>>(n, d_ian, d_jan, d_rowptr, d_col, d_val,
                                                                 const_cast<double *>(d_a));
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

}  // namespace rs
