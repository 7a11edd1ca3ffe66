/*
 * rootsolve_b200_model.cuh -- how a compiled model plugs into the GPU solver.
 *
 * rootSolve's compiled models are C/Fortran functions in a DLL with the
 * signature (src/steady.h:34, vignettes/rootSolve.Rnw:888-934)
 *
 *     void func(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
 *
 * plus an optional initialiser  void initfunc(void (*steadyparms)(int *, double *))
 * that receives the parameter vector (src/steady.h:35, inst/dynload/anoxmodc.c:21-25).
 * Here the same two functions are written once in a .cu file:
 *
 *     #include "rootsolve_b200_model.cuh"
 *     RS_MODEL_PARMS(3)                       // double parms[3] in __constant__ memory
 *     __device__ void mymod(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
 *     {
 *         RS_FOR_EACH(i, *neq) ydot[i] = -parms[0] * y[i];    // each thread does its share
 *     }
 *     RS_MODEL(mymod, 3)                      // exports  rs_model_mymod()  from the model .so
 *
 * Contract of the device function (what differs from the CPU DLL):
 *   * it is executed by EVERY thread of a (blocks x batch) grid, all with the
 *     same six arguments; it must compute ydot[i] for the indices it owns --
 *     RS_FOR_EACH / RS_THREAD_ID / RS_THREAD_COUNT partition 0..n-1 over the
 *     threads of one evaluation -- and must not write y;
 *   * y and ydot always use GLOBAL indices (the layout of the R state vector),
 *     also when the solve is slab-partitioned over several GPUs: a rank then
 *     owns planes [RS_SLAB_J0(ip), RS_SLAB_J1(ip)) of the slowest grid index
 *     and must compute ydot only there (its y holds valid values on those
 *     planes and one halo plane on either side);
 *   * ip[] and yout[] have the reference's layout (src/steady_utils.c:109-124):
 *     ip[0]=nout, ip[1]=length(yout), ip[2]=length(ip)=3+length(ipar), ip[3..]=ipar;
 *     yout[0..nout) outputs, then rpar.  The two ints after the user's ipar are
 *     the slab bounds (J1 < 0: not partitioned).
 *
 * Accessor form (optional, faster).  The finite-difference Jacobian evaluates the model on ngp perturbed copies of
 * the state (src/sparse.f:497-508).  A model written as above needs those copies in memory.  A model that takes its
 * state as an `rs_state' instead of a `double *' --
 *
 *     __device__ void mymod(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
 *     {   RS_FOR_EACH(i, *neq) ydot[i] = -parms[0] * y[i];   }        // y[i], y + offset work as on a pointer
 *     RS_MODEL_ACC(mymod, 3)
 *
 * -- reads the ONE state vector of the solve, and y[i] adds the solver's increment (Perturb, src/steady.f:208-226)
 * on the fly when unknown i belongs to the column group the evaluation perturbs: the ngp copies are never written
 * or read.  The body of the function does not change.  RS_FOR_EACH_CELL walks a box of grid cells line by line
 * without an integer division per cell.
 *
 * Row form (optional, on top of either form above).  xSparseJacob (src/sparse.f:461-525) perturbs all columns of a
 * group at once because a model call is expensive and computes every row; what it extracts for entry (i, j) is
 * f_i with column j alone perturbed -- the other columns of the group do not reach row i, that is what a group is.
 * A model that can evaluate ONE row,
 *
 *     __device__ double mymod_row(long long i, double t, rs_state y, const double *yout, const int *ip)
 *     {   return -parms[0] * y[i];   }           // must be, bit for bit, what mymod writes to ydot[i]
 *     RS_MODEL_ACC_ROW(mymod, mymod_row, 3)
 *
 * lets the solver compute every entry of the Jacobian from registers: per row one evaluation of the row function for
 * f_i(y) and one per entry of the row, the state read once, the values written once, no residual vectors in between.
 * The entries are bit-identical to the grouped evaluation's.
 *
 * The kernel that calls func is instantiated in the model's own translation
 * unit (template argument = the function), so the residual is inlined into
 * the batched launch; the solver library reaches it through the host function
 * pointer in rs_model_desc.  This is the GPU analogue of R's
 * getNativeSymbolInfo(func, dllname)$address (R/stodes.R:212-220).
 */
#ifndef ROOTSOLVE_B200_MODEL_CUH
#define ROOTSOLVE_B200_MODEL_CUH

#include <cuda_runtime.h>

#define RS_MODEL_ABI 2

typedef void rs_deriv_func_t(int *, double *, double *, double *, double *, int *);
typedef void rs_init_func_t(void (*)(int *, double *));

struct rs_model_desc {
    int abi;
    const char *name;
    int nparms; /* length the initialiser asks for; 0 = no initialiser */
    /* Evaluates nbatch residuals in ONE launch: evaluation b reads y + b*ld, writes ydot + b*ld and
     * uses yout + b*ldout.  work = number of unknowns this rank owns (sizes the grid). */
    int (*launch)(void *stream, int neq, double t, const double *y, double *ydot, long long ld, int nbatch,
                  double *yout, long long ldout, int *ip, long long work);
    rs_init_func_t *initfunc; /* reference signature; copies parms to the device */
    rs_init_func_t *initforc;
    /* Accessor-form models only (else NULL): evaluates the model on y with the columns of group g0 + b perturbed,
     * b = 0 .. ngroups-1, in ONE launch; evaluation b writes ydot + b*ld and uses yout + b*ldout.
     * group_of[i] = 0-based column group of unknown i. */
    int (*launch_groups)(void *stream, int neq, double t, const double *y, const int *group_of, int g0, int ngroups,
                         double *ydot, long long ld, double *yout, long long ldout, int *ip, long long work);
    /* Models that also export a ROW function (else NULL): the whole finite-difference Jacobian of rows [row0, row1) in
     * one launch, val[p] = (f_i(y + delta_j e_j) - f_i(y)) / ((y_j + delta_j) - y_j) for entry p = (i, j = col[p]) of the
     * row-wise arrays; nothing but y, the pattern and val touches memory. */
    int (*launch_fd_rows)(void *stream, int row0, int row1, double t, const double *y, const int *rowptr, const int *col,
                          double *val, const double *yout, const int *ip);
};

/* The state vector as an accessor-form model sees it. */
struct rs_state {
    const double *y;
    const int *grp; /* column group of every unknown; may be NULL when g < 0 */
    int g;          /* group whose columns carry the finite-difference increment; < 0: none */
    long long pj;   /* single unknown that carries the increment (row form); < 0: none */
    __device__ __forceinline__ double operator[](long long i) const
    {
        const double v = y[i];
        if ((g >= 0 && grp[i] == g) || i == pj) {
            double d = fabs(v) * 1.0e-8; /* Perturb, src/steady.f:219-224 */
            if (d < 1.0e-8) d = 1.0e-8;
            return v + d;
        }
        return v;
    }
    __device__ __forceinline__ rs_state operator+(long long off) const
    {
        rs_state s = {y + off, grp ? grp + off : grp, g, pj >= 0 ? pj - off : pj};
        return s;
    }
};
typedef void rs_deriv_acc_func_t(int *, double *, rs_state, double *, double *, int *);

/* Line-by-line walk over the cells (j, k, l) of planes [j0, j1) of a box with extents (*, d2, d3), l fastest: the
 * threads of a block share lines of l, a thread finds its (j, k) by one division when it starts and by additions
 * afterwards.  A 2-D grid (a, b) is walked as planes [j0, j1) of (*, 1, b).
 *     RS_FOR_EACH_CELL(it, l, j0, j1, d2, d3) { ... it.j, it.k, l ... }                                          */
struct rs_lines {
    int j, k, ln, nl, stride, d2, tl, lstep;
    __device__ __forceinline__ rs_lines(int j0, int j1, int d2_, int d3)
    {
        int sh = 0;
        while ((1 << sh) < d3 && (1 << sh) < (int)blockDim.x) sh++;
        lstep = 1 << sh;
        tl = threadIdx.x & (lstep - 1);
        const int per_block = (int)blockDim.x >> sh;
        d2 = d2_;
        nl = (j1 - j0) * d2_;
        ln = blockIdx.x * per_block + ((int)threadIdx.x >> sh);
        stride = gridDim.x * per_block;
        j = j0 + ln / d2_;
        k = ln - (j - j0) * d2_;
    }
    __device__ __forceinline__ bool ok() const { return ln < nl; }
    __device__ __forceinline__ void next()
    {
        ln += stride;
        k += stride;
        while (k >= d2) { k -= d2; j++; }
    }
};
#define RS_FOR_EACH_CELL(it, l, j0, j1, d2, d3) \
    for (rs_lines it((j0), (j1), (d2), (d3)); it.ok(); it.next()) \
        for (int l = it.tl; l < (d3); l += it.lstep)

#define RS_THREAD_ID ((long long)blockIdx.x * blockDim.x + threadIdx.x)
#define RS_THREAD_COUNT ((long long)gridDim.x * blockDim.x)
#define RS_FOR_EACH(i, n) for (long long i = RS_THREAD_ID; i < (long long)(n); i += RS_THREAD_COUNT)
#define RS_SLAB_J0(ip) ((ip)[(ip)[2]])
#define RS_SLAB_J1(ip) ((ip)[(ip)[2] + 1])

template <rs_deriv_func_t F>
__global__ void __launch_bounds__(256) rs_residual_batched_kernel(int neq, double t, const double *y,
                                                                   double *ydot, long long ld, double *yout,
                                                                   long long ldout, int *ip)
{
    const long long b = blockIdx.y;
    int n = neq;
    double tt = t;
    F(&n, &tt, const_cast<double *>(y) + b * ld, ydot + b * ld, yout + b * ldout, ip);
}

template <rs_deriv_func_t F>
static int rs_launch_batched(void *stream, int neq, double t, const double *y, double *ydot, long long ld,
                             int nbatch, double *yout, long long ldout, int *ip, long long work)
{
    if (work <= 0 || nbatch <= 0) return 0;
    const int threads = 256;
    long long blocks = (work + threads - 1) / threads;
    /* One wave: the whole (blocks x nbatch) grid is resident at once -- SMs x the CTAs of THIS model's kernel an SM
     * holds (its register count decides) -- and every thread strides over its share; a grid that is a fraction of
     * a wave larger leaves the SMs idle for most of a second wave. */
    static int resident = 0;
    if (resident == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rs_residual_batched_kernel<F>, threads, 0) == cudaSuccess &&
            sms > 0 && per_sm > 0)
            resident = sms * per_sm;
        else
            resident = 148 * 4;
    }
    long long cap = resident / nbatch;
    if (cap < 1) cap = 1;
    if (blocks > cap) blocks = cap;
    dim3 grid((unsigned)blocks, (unsigned)nbatch);
    rs_residual_batched_kernel<F><<<grid, threads, 0, (cudaStream_t)stream>>>(neq, t, y, ydot, ld, yout, ldout, ip);
    return (int)cudaGetLastError();
}

/* ---- accessor form ---- */
template <rs_deriv_acc_func_t F>
__global__ void __launch_bounds__(256) rs_residual_batched_acc_kernel(int neq, double t, const double *y, double *ydot,
                                                                       long long ld, double *yout, long long ldout, int *ip)
{
    const long long b = blockIdx.y;
    int n = neq;
    double tt = t;
    rs_state s = {y + b * ld, nullptr, -1, -1};
    F(&n, &tt, s, ydot + b * ld, yout + b * ldout, ip);
}

template <rs_deriv_acc_func_t F>
__global__ void __launch_bounds__(256) rs_residual_groups_kernel(int neq, double t, const double *y, const int *group_of, int g0,
                                                                  double *ydot, long long ld, double *yout, long long ldout,
                                                                  int *ip)
{
    const long long b = blockIdx.y;
    int n = neq;
    double tt = t;
    rs_state s = {y, group_of, g0 + (int)b, -1};
    F(&n, &tt, s, ydot + b * ld, yout + b * ldout, ip);
}

template <typename K>
static long long rs_wave_blocks_(K kernel, int threads, int nbatch, int *cache)
{
    if (*cache == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) == cudaSuccess && sms > 0 && per_sm > 0)
            *cache = sms * per_sm;
        else
            *cache = 148 * 4;
    }
    long long cap = *cache / nbatch;
    return cap < 1 ? 1 : cap;
}

template <rs_deriv_acc_func_t F>
static int rs_launch_batched_acc(void *stream, int neq, double t, const double *y, double *ydot, long long ld,
                                 int nbatch, double *yout, long long ldout, int *ip, long long work)
{
    if (work <= 0 || nbatch <= 0) return 0;
    const int threads = 256;
    static int resident = 0;
    long long blocks = (work + threads - 1) / threads;
    const long long cap = rs_wave_blocks_(rs_residual_batched_acc_kernel<F>, threads, nbatch, &resident);
    if (blocks > cap) blocks = cap;
    dim3 grid((unsigned)blocks, (unsigned)nbatch);
    rs_residual_batched_acc_kernel<F><<<grid, threads, 0, (cudaStream_t)stream>>>(neq, t, y, ydot, ld, yout, ldout, ip);
    return (int)cudaGetLastError();
}

template <rs_deriv_acc_func_t F>
static int rs_launch_groups(void *stream, int neq, double t, const double *y, const int *group_of, int g0, int ngroups,
                            double *ydot, long long ld, double *yout, long long ldout, int *ip, long long work)
{
    if (work <= 0 || ngroups <= 0) return 0;
    const int threads = 256;
    static int resident = 0;
    long long blocks = (work + threads - 1) / threads;
    /* One resident wave over all evaluations (grid.y = evaluation), every thread striding over cells: the evaluations
     * sweep the state together and share its lines in L2.  (Looping over the evaluations inside one kernel whose grid
     * covers the cells once was measured 4.5 times slower on B200 at 256^3: 13 widely spaced output streams per block.) */
    const long long cap = rs_wave_blocks_(rs_residual_groups_kernel<F>, threads, ngroups, &resident);
    if (blocks > cap) blocks = cap;
    dim3 grid((unsigned)blocks, (unsigned)ngroups);
    rs_residual_groups_kernel<F><<<grid, threads, 0, (cudaStream_t)stream>>>(neq, t, y, group_of, g0, ydot, ld, yout, ldout, ip);
    return (int)cudaGetLastError();
}

/* ---- row form: the fused finite-difference Jacobian ---- */
typedef double rs_row_func_t(long long, double, rs_state, const double *, const int *);
#define RS_FD_TILE 256
#define RS_FD_STAGE (RS_FD_TILE * 9)

template <rs_row_func_t F>
__global__ void __launch_bounds__(RS_FD_TILE) rs_fd_rows_kernel(int row0, int row1, double t, const double *__restrict__ y,
                                                                 const int *__restrict__ rowptr, const int *__restrict__ col,
                                                                 double *__restrict__ val, const double *yout, const int *ip)
{
    /* a tile of 256 consecutive rows owns one contiguous slice of val[]: staged in shared memory, written coalesced */
    __shared__ double sv[RS_FD_STAGE];
    __shared__ int s_base, s_end;
    const int ntiles = (row1 - row0 + RS_FD_TILE - 1) / RS_FD_TILE;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int i0 = row0 + tile * RS_FD_TILE, i1 = min(row1, i0 + RS_FD_TILE);
        const int i = i0 + (int)threadIdx.x;
        const bool valid = i < i1;
        if (threadIdx.x == 0) { s_base = rowptr[i0]; s_end = rowptr[i1]; }
        __syncthreads();
        const int base_p = s_base, cnt = s_end - s_base;
        const bool staged = cnt <= RS_FD_STAGE;
        if (valid) {
            const rs_state s0 = {y, nullptr, -1, -1};
            const double f0 = F(i, t, s0, yout, ip);
            const int pe = rowptr[i + 1];
            for (int p = rowptr[i]; p < pe; p++) {
                const int j = col[p];
                const double yj = y[j];
                double d = fabs(yj) * 1.0e-8; /* Perturb, src/steady.f:219-224 */
                if (d < 1.0e-8) d = 1.0e-8;
                const double divdelt = (yj + d) - yj; /* the rounded difference, src/sparse.f:511 */
                const rs_state sj = {y, nullptr, -1, j};
                const double a = (F(i, t, sj, yout, ip) + (-f0)) / divdelt;
                if (staged) sv[p - base_p] = a; else val[p] = a;
            }
        }
        __syncthreads();
        if (staged)
            for (int e = threadIdx.x; e < cnt; e += RS_FD_TILE) val[base_p + e] = sv[e];
        __syncthreads();
    }
}

template <rs_row_func_t F>
static int rs_launch_fd_rows(void *stream, int row0, int row1, double t, const double *y, const int *rowptr, const int *col,
                             double *val, const double *yout, const int *ip)
{
    if (row1 <= row0) return 0;
    long long blocks = ((long long)(row1 - row0) + RS_FD_TILE - 1) / RS_FD_TILE;
    if (blocks > 148 * 16) blocks = 148 * 16;
    rs_fd_rows_kernel<F><<<(unsigned)blocks, RS_FD_TILE, 0, (cudaStream_t)stream>>>(row0, row1, t, y, rowptr, col, val, yout, ip);
    return (int)cudaGetLastError();
}

/* N parameters in __constant__ memory + the reference-style initialiser that fills them */
#define RS_MODEL_PARMS(N)                                                        \
    static __constant__ double parms[N];                                         \
    static double rs_parms_host_[N];                                             \
    static void rs_model_init_(void (*steadyparms)(int *, double *))             \
    {                                                                            \
        int n_ = N;                                                              \
        steadyparms(&n_, rs_parms_host_);                                        \
        cudaMemcpyToSymbol(parms, rs_parms_host_, sizeof(double) * N);           \
    }

/* N forcing function values in __constant__ memory + the initialiser `initforc` of the reference's compiled
 * models (deSolve's convention; R/stodes.R:222-237, src/steady_utils.c:62-77).  A steady-state solve sees the
 * forcings at ONE time, so the values are constants of the solve. */
#define RS_MODEL_FORCS(N)                                                        \
    static __constant__ double forcs[N];                                         \
    static double rs_forcs_host_[N];                                             \
    static void rs_model_initforc_(void (*steadyforcs)(int *, double *))         \
    {                                                                            \
        int n_ = N;                                                              \
        steadyforcs(&n_, rs_forcs_host_);                                        \
        cudaMemcpyToSymbol(forcs, rs_forcs_host_, sizeof(double) * N);           \
    }

#define RS_MODEL_NAMED_F(NAME, FUNC, NPARMS, INITFN, FORCFN)                     \
    extern "C" const rs_model_desc *rs_model_##NAME(void)                        \
    {                                                                            \
        static const rs_model_desc d = {RS_MODEL_ABI, #NAME, NPARMS, rs_launch_batched<FUNC>, INITFN, FORCFN, nullptr, nullptr}; \
        return &d;                                                               \
    }
#define RS_MODEL_NAMED(NAME, FUNC, NPARMS, INITFN) RS_MODEL_NAMED_F(NAME, FUNC, NPARMS, INITFN, nullptr)

#define RS_MODEL_ACC_NAMED_F(NAME, FUNC, NPARMS, INITFN, FORCFN)                 \
    extern "C" const rs_model_desc *rs_model_##NAME(void)                        \
    {                                                                            \
        static const rs_model_desc d = {RS_MODEL_ABI, #NAME, NPARMS, rs_launch_batched_acc<FUNC>, INITFN, FORCFN, \
                                        rs_launch_groups<FUNC>, nullptr};        \
        return &d;                                                               \
    }
#define RS_MODEL_ACC_ROW_NAMED_F(NAME, FUNC, ROWFUNC, NPARMS, INITFN, FORCFN)    \
    extern "C" const rs_model_desc *rs_model_##NAME(void)                        \
    {                                                                            \
        static const rs_model_desc d = {RS_MODEL_ABI, #NAME, NPARMS, rs_launch_batched_acc<FUNC>, INITFN, FORCFN, \
                                        rs_launch_groups<FUNC>, rs_launch_fd_rows<ROWFUNC>}; \
        return &d;                                                               \
    }
/* accessor-form model that also exports its row function */
#define RS_MODEL_ACC_ROW(FUNC, ROWFUNC, NPARMS) RS_MODEL_ACC_ROW_NAMED_F(FUNC, FUNC, ROWFUNC, NPARMS, rs_model_init_, nullptr)
/* accessor-form model with RS_MODEL_PARMS(NPARMS) declared earlier in the same file */
#define RS_MODEL_ACC(FUNC, NPARMS) RS_MODEL_ACC_NAMED_F(FUNC, FUNC, NPARMS, rs_model_init_, nullptr)

/* model with RS_MODEL_PARMS(NPARMS) declared earlier in the same file */
#define RS_MODEL(FUNC, NPARMS) RS_MODEL_NAMED(FUNC, FUNC, NPARMS, rs_model_init_)
/* model with RS_MODEL_PARMS(NPARMS) and RS_MODEL_FORCS(...) declared earlier in the same file */
#define RS_MODEL_WITH_FORCS(FUNC, NPARMS) RS_MODEL_NAMED_F(FUNC, FUNC, NPARMS, rs_model_init_, rs_model_initforc_)
/* model without parameters */
#define RS_MODEL_NOPARMS(FUNC) RS_MODEL_NAMED(FUNC, FUNC, 0, nullptr)

#endif /* ROOTSOLVE_B200_MODEL_CUH */
