// structure.cu -- sparsity structure of the Jacobian on the device.
//
// K0: ian/jan for 1-D / 2-D / 3-D multi-species grids, bit-exact with the
//     reference's sparse1d / sparse2d / sparse3d (src/sparsecoeff.f:5-312): one
//     thread per column emits its row indices in the reference's order (a count
//     pass, a device prefix sum, a fill pass).
// CSR: the reference's (ian, jan) lists the rows of every COLUMN of J
//     (src/sparse.f:309-316); the Krylov solver needs J by rows, so the
//     pattern is transposed once per solve (histogram, prefix sum, fill, per-row
//     sort) and the values are assembled straight into the row-wise arrays.
#include "rs_internal.cuh"
#include "solver.cuh"
#include "grid_rows.cuh"

namespace rs {

// ---- exclusive prefix sum (int32) ------------------------------------------
// Three-phase scan: each 512-thread block scans 2048 items in shared memory and
// emits its total; the totals are scanned recursively; a last pass adds the
// block offsets.  out[i] = base + sum(in[0..i)), and out[n] = base + total when
// write_total is set (out then has n+1 entries).
constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 4;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_kernel(const int *__restrict__ in, int *__restrict__ out,
                                                                 int *__restrict__ block_sums, int n)
{
    __shared__ int warp_tot[SCAN_THREADS / 32];
    const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    int v[SCAN_ITEMS], tsum = 0;
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS; q++) {
        v[q] = (base + q < n) ? in[base + q] : 0;
        tsum += v[q];
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int w = (lane < SCAN_THREADS / 32) ? warp_tot[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        if (lane < SCAN_THREADS / 32) warp_tot[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    int run = incl - tsum + (wid > 0 ? warp_tot[wid - 1] : 0);
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS; q++) {
        if (base + q < n) out[base + q] = run;
        run += v[q];
    }
    if (threadIdx.x == SCAN_THREADS - 1) block_sums[blockIdx.x] = run;
}

__global__ void scan_add_kernel(int *__restrict__ out, const int *__restrict__ block_off, int n, int base,
                                int write_total, const int *__restrict__ last_in)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] += block_off[i / SCAN_TILE] + base;
    if (write_total && i == n - 1) {
        // out[n] = out[n-1] + in[n-1]; out[n-1] is final after this thread's own update
        out[n] = out[n - 1] + last_in[n - 1];
    }
}

int exclusive_scan(Ctx &c, const int *in, int *out, int n, int base, bool write_total)
{
    if (n <= 0) {
        if (write_total) RS_CUDA(cudaMemcpyAsync(out, &base, sizeof(int), cudaMemcpyHostToDevice, c.stream));
        return RS_OK;
    }
    const int nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    DevBuf<int> sums, offs;
    RS_TRY(sums.alloc(nb));
    RS_TRY(offs.alloc(nb + 1));
    {
        Launch L(c, K_SCAN, 8.0 * n);
        scan_tile_kernel<<<nb, SCAN_THREADS, 0, c.stream>>>(in, out, sums.p, n);
    }
    if (nb > 1) {
        RS_TRY(exclusive_scan(c, sums.p, offs.p, nb, 0, false));
    } else {
        RS_CUDA(cudaMemsetAsync(offs.p, 0, sizeof(int), c.stream));
    }
    {
        Launch L(c, K_SCAN, 8.0 * n);
        scan_add_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(out, offs.p, n, base, write_total ? 1 : 0, in);
    }
    RS_CUDA(cudaGetLastError());
    RS_CUDA(cudaStreamSynchronize(c.stream));  // sums/offs are released on return
    return RS_OK;
}

// ---- K0: grid structures (emit_column lives in grid_rows.cuh) ----------------------
__global__ void structure_count_kernel(GridDesc g, int n, int *__restrict__ counts)
{
    const int M = blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (M > n) return;
    int c = 0;
    emit_column(g, M, [&](int) { c++; });
    counts[M - 1] = c;
}

__global__ void structure_fill_kernel(GridDesc g, int n, const int *__restrict__ ian, int *__restrict__ jan)
{
    const int M = blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (M > n) return;
    int pos = ian[M - 1] - 1;
    emit_column(g, M, [&](int row) { jan[pos++] = row; });
}

// Builds ian (n+1, 1-based) and jan (1-based rows) on the device.  d_jan is
// allocated here once the count is known.
int build_grid_structure(Ctx &c, const GridDesc &g, int n, DevBuf<int> &d_ian, DevBuf<int> &d_jan, int *nnz)
{
    if (g.nspec <= 0 || g.ncell <= 0 || (long long)g.nspec * g.ncell != n) {
        set_error("cannot generate sparse jacobian - N and nspec not compatible");
        return RS_ESTRUCT;
    }
    DevBuf<int> counts;
    RS_TRY(counts.alloc(n));
    RS_TRY(d_ian.alloc((size_t)n + 1));
    {
        Launch L(c, K_STRUCT, 4.0 * n);
        structure_count_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(g, n, counts.p);
    }
    RS_TRY(exclusive_scan(c, counts.p, d_ian.p, n, 1, true));
    int last = 0;
    RS_CUDA(cudaMemcpyAsync(&last, d_ian.p + n, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    *nnz = last - 1;
    RS_TRY(d_jan.alloc((size_t)*nnz));
    {
        Launch L(c, K_STRUCT, 4.0 * n + 4.0 * *nnz);
        structure_fill_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(g, n, d_ian.p, d_jan.p);
    }
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

// ---- mapped (masked) grids: sparse2dmap / sparse3dmap, src/testMap.f:25-227 ----------
// pres[M-1] = new number (1..n, increasing with M) of full-grid state M, <= 0 if absent.  The reference walks the
// full grid and appends to column pres(M) the NEW numbers of the present neighbours, in sparse2d/sparse3d order.
// Here: invert the map (inv[c] = M), then one thread per present column runs emit_column on its full-grid
// position and keeps the present rows (count pass, prefix sum, fill pass as above).
__global__ void map_invert_kernel(const int *__restrict__ pres, int totn, int n, int *__restrict__ inv, int *__restrict__ flag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int p = (i < totn) ? pres[i] : 0;
    if (p > n) atomicMax(&flag[0], 1);
    else if (p > 0) inv[p - 1] = i + 1;
    const unsigned present = __ballot_sync(0xffffffffu, p > 0);
    if ((threadIdx.x & 31) == 0 && present) atomicAdd(&flag[1], __popc(present));
}

__global__ void map_check_kernel(const int *__restrict__ inv, int n, int *__restrict__ flag)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    if (inv[c] == 0 || (c + 1 < n && inv[c] >= inv[c + 1])) atomicMax(&flag[0], 2);
}

template <bool FILL>
__global__ void structure_map_kernel(GridDesc g, int n, int totn, const int *__restrict__ pres, const int *__restrict__ inv,
                                     const int *__restrict__ ian, int *__restrict__ out)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    int cnt = 0, pos = FILL ? ian[c] - 1 : 0;
    emit_column(g, inv[c], [&](int row) {
        // a row outside 1..totn can only come from the 2-D wrap of src/testMap.f:97 when d1 > d2 (the reference
        // reads beyond `pres` there); it counts as absent
        const int p = (row >= 1 && row <= totn) ? pres[row - 1] : 0;
        if (p > 0) {
            if (FILL) out[pos++] = p; else cnt++;
        }
    });
    if (!FILL) out[c] = cnt;
}

// g describes the FULL grid (g.nspec * g.ncell = totn = length of h_pres); n = number of present states.
int build_mapped_structure(Ctx &c, const GridDesc &g, int n, int totn, const int *h_pres, DevBuf<int> &d_ian,
                           DevBuf<int> &d_jan, int *nnz)
{
    if (g.nspec <= 0 || (n / g.nspec) * g.nspec != n) {           // src/testMap.f:51-54, 141-144
        set_error("cannot generate sparse jacobian - N and nspec not compatible");
        return RS_ESTRUCT;
    }
    if (g.ncell <= 0 || (long long)g.nspec * g.ncell != totn) {
        set_error("mapped grid: the map has %d entries, nspec*prod(dimens) = %lld", totn, (long long)g.nspec * g.ncell);
        return RS_EINVAL;
    }
    DevBuf<int> pres, inv, flag, counts;
    RS_TRY(pres.alloc(totn)); RS_TRY(inv.alloc(n)); RS_TRY(flag.alloc(2)); RS_TRY(counts.alloc(n));
    RS_TRY(pres.upload(h_pres, totn, c.stream));
    RS_TRY(inv.zero(c.stream));
    RS_TRY(flag.zero(c.stream));
    {
        Launch L(c, K_STRUCT, 4.0 * totn + 4.0 * n);
        map_invert_kernel<<<(totn + 255) / 256, 256, 0, c.stream>>>(pres.p, totn, n, inv.p, flag.p);
    }
    {
        Launch L(c, K_STRUCT, 4.0 * n);
        map_check_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(inv.p, n, flag.p);
    }
    int hflag[2] = {0, 0};
    RS_CUDA(cudaMemcpyAsync(hflag, flag.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    if (hflag[0] || hflag[1] != n) {
        set_error("mapped grid: the map must number the %d present states 1..%d in grid order (%d entries are present)", n, n,
                  hflag[1]);
        return RS_EINVAL;
    }
    RS_TRY(d_ian.alloc((size_t)n + 1));
    {
        Launch L(c, K_STRUCT, 8.0 * n);
        structure_map_kernel<false><<<(n + 255) / 256, 256, 0, c.stream>>>(g, n, totn, pres.p, inv.p, nullptr, counts.p);
    }
    RS_TRY(exclusive_scan(c, counts.p, d_ian.p, n, 1, true));
    int last = 0;
    RS_CUDA(cudaMemcpyAsync(&last, d_ian.p + n, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    *nnz = last - 1;
    RS_TRY(d_jan.alloc((size_t)*nnz));
    {
        Launch L(c, K_STRUCT, 8.0 * n + 4.0 * *nnz);
        structure_map_kernel<true><<<(n + 255) / 256, 256, 0, c.stream>>>(g, n, totn, pres.p, inv.p, d_ian.p, d_jan.p);
    }
    RS_CUDA(cudaGetLastError());
    RS_CUDA(cudaStreamSynchronize(c.stream));  // pres/inv are released on return
    return RS_OK;
}

// ---- CSC -> CSR pattern ------------------------------------------------------
__global__ void row_hist_kernel(const int *__restrict__ jan, int nnz, int *__restrict__ cnt)
{
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (long long)gridDim.x * blockDim.x)
        atomicAdd(&cnt[jan[k] - 1], 1);
}

__global__ void row_fill_kernel(const int *__restrict__ ian, const int *__restrict__ jan, int n,
                                const int *__restrict__ rowptr, int *__restrict__ cursor, int *__restrict__ col)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;  // column, 0-based
    if (j >= n) return;
    for (int k = ian[j] - 1; k < ian[j + 1] - 1; k++) {
        const int i = jan[k] - 1;
        const int pos = atomicAdd(&cursor[i], 1);
        col[rowptr[i] + pos] = j;
    }
}

// sorts every row's column indices ascending; flags duplicates
__global__ void row_sort_kernel(const int *__restrict__ rowptr, int n, int *__restrict__ col, int *__restrict__ dup_flag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b = rowptr[i], e = rowptr[i + 1];
    for (int p = b + 1; p < e; p++) {
        const int key = col[p];
        int q = p - 1;
        while (q >= b && col[q] > key) { col[q + 1] = col[q]; q--; }
        col[q + 1] = key;
    }
    for (int p = b + 1; p < e; p++)
        if (col[p] == col[p - 1]) atomicMax(dup_flag, i + 1);
}

// CSR pattern of J (0-based rowptr/col) from the column-wise (ian, jan).
int build_csr_pattern(Ctx &c, int n, int nnz, const int *d_ian, const int *d_jan, DevBuf<int> &rowptr,
                      DevBuf<int> &col)
{
    DevBuf<int> cnt, flag;
    RS_TRY(cnt.alloc(n));
    RS_TRY(flag.alloc(1));
    RS_TRY(rowptr.alloc((size_t)n + 1));
    RS_TRY(col.alloc((size_t)nnz));
    RS_TRY(cnt.zero(c.stream));
    RS_TRY(flag.zero(c.stream));
    {
        Launch L(c, K_TRANSPOSE, 8.0 * nnz);
        row_hist_kernel<<<grid_for(nnz, 256, 148 * 16), 256, 0, c.stream>>>(d_jan, nnz, cnt.p);
    }
    RS_TRY(exclusive_scan(c, cnt.p, rowptr.p, n, 0, true));
    RS_TRY(cnt.zero(c.stream));
    {
        Launch L(c, K_TRANSPOSE, 12.0 * nnz);
        row_fill_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(d_ian, d_jan, n, rowptr.p, cnt.p, col.p);
    }
    {
        Launch L(c, K_TRANSPOSE, 8.0 * nnz);
        row_sort_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(rowptr.p, n, col.p, flag.p);
    }
    int dup = 0;
    RS_CUDA(cudaMemcpyAsync(&dup, flag.p, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    if (dup) {
        set_error("sparse solver: duplicate entry in a (row nr: %d)", dup);
        return RS_EDUPLICATE;
    }
    return RS_OK;
}

// ---- row-wise pattern of a non-cyclic grid in closed form ------------------------------------------------------
// Without wrap-around the pattern of sparse1d / 2d / 3d (src/sparsecoeff.f:5-312) is structurally symmetric: row
// (species s, cell) holds the other species of the cell and the cell's grid neighbours.  Ascending column order:
// the species before s, then -j -k -l, the diagonal, +l +k +j of species s, then the species after s.  Writing that
// down directly replaces the histogram, the scattered fill and the per-row sort of the general transposition
// (1.9 -> 0.3 ms of set-up at 256^3).  RS_CHECK_GRID_PATTERN=1 builds both and compares them (tests).
struct GridBox { int nspec, ncell, b1, b2, b3; };

__device__ __forceinline__ int grid_row_emit(const GridBox &g, int r, int *out)
{
    const int s = r / g.ncell, cell = r - s * g.ncell;
    const int s1 = g.b2 * g.b3;
    const int j = cell / s1, rem = cell - j * s1, k = rem / g.b3, l = rem - k * g.b3;
    const int base = s * g.ncell;
    int m = 0;
    for (int q = 0; q < s; q++) { if (out) out[m] = q * g.ncell + cell; m++; }
    if (j > 0) { if (out) out[m] = base + cell - s1; m++; }
    if (k > 0) { if (out) out[m] = base + cell - g.b3; m++; }
    if (l > 0) { if (out) out[m] = base + cell - 1; m++; }
    if (out) out[m] = r;
    m++;
    if (l < g.b3 - 1) { if (out) out[m] = base + cell + 1; m++; }
    if (k < g.b2 - 1) { if (out) out[m] = base + cell + g.b3; m++; }
    if (j < g.b1 - 1) { if (out) out[m] = base + cell + s1; m++; }
    for (int q = s + 1; q < g.nspec; q++) { if (out) out[m] = q * g.ncell + cell; m++; }
    return m;
}

__global__ void grid_row_count_kernel(GridBox g, int n, int *__restrict__ cnt)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) cnt[r] = grid_row_emit(g, r, nullptr);
}

__global__ void grid_row_fill_kernel(GridBox g, int n, const int *__restrict__ rowptr, int *__restrict__ col)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) grid_row_emit(g, r, col + rowptr[r]);
}

__global__ void int_mismatch_kernel(long long n, const int *__restrict__ a, const int *__restrict__ b, int *__restrict__ count)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        if (a[i] != b[i]) atomicAdd(count, 1);
}

bool grid_pattern_applicable(const GridDesc &g)
{
    if (g.type != RS_TYPE_1D && g.type != RS_TYPE_2D && g.type != RS_TYPE_3D) return false;
    return !g.cyc[0] && !g.cyc[1] && !g.cyc[2] && !getenv("RS_NO_GRID_PATTERN");
}

int build_csr_pattern_grid(Ctx &c, const GridDesc &g, int n, int nnz, const int *d_ian, const int *d_jan, DevBuf<int> &rowptr,
                           DevBuf<int> &col)
{
    GridBox b;
    b.nspec = g.nspec; b.ncell = g.ncell;
    const int nd = g.type == RS_TYPE_3D ? 3 : g.type == RS_TYPE_2D ? 2 : 1;
    b.b1 = nd == 3 ? g.d[0] : 1;
    b.b2 = nd == 3 ? g.d[1] : nd == 2 ? g.d[0] : 1;
    b.b3 = g.d[nd - 1];
    DevBuf<int> cnt;
    RS_TRY(cnt.alloc(n));
    RS_TRY(rowptr.alloc((size_t)n + 1));
    RS_TRY(col.alloc((size_t)nnz));
    {
        Launch L(c, K_TRANSPOSE, 4.0 * n);
        grid_row_count_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(b, n, cnt.p);
    }
    RS_TRY(exclusive_scan(c, cnt.p, rowptr.p, n, 0, true));
    int total = -1;
    RS_CUDA(cudaMemcpyAsync(&total, rowptr.p + n, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    if (total != nnz) return build_csr_pattern(c, n, nnz, d_ian, d_jan, rowptr, col);   // not the stencil the closed form assumes
    {
        Launch L(c, K_TRANSPOSE, 4.0 * n + 4.0 * nnz);
        grid_row_fill_kernel<<<(n + 255) / 256, 256, 0, c.stream>>>(b, n, rowptr.p, col.p);
    }
    RS_CUDA(cudaGetLastError());
    if (getenv("RS_CHECK_GRID_PATTERN")) {
        DevBuf<int> rp2, col2, bad;
        RS_TRY(build_csr_pattern(c, n, nnz, d_ian, d_jan, rp2, col2));
        RS_TRY(bad.alloc(1));
        RS_TRY(bad.zero(c.stream));
        int_mismatch_kernel<<<grid_for((long long)n + 1, 256), 256, 0, c.stream>>>((long long)n + 1, rowptr.p, rp2.p, bad.p);
        int_mismatch_kernel<<<grid_for(nnz, 256), 256, 0, c.stream>>>(nnz, col.p, col2.p, bad.p);
        int nbad = 0;
        RS_CUDA(cudaMemcpyAsync(&nbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
        RS_CUDA(cudaStreamSynchronize(c.stream));
        if (nbad) {
            set_error("internal: the closed-form row pattern of the grid differs from the transposed structure in %d places", nbad);
            return RS_ESTRUCT;
        }
    }
    return RS_OK;
}

}  // namespace rs
