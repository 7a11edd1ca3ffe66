// api.cu -- the C ABI (include/rootsolve_b200.h) and the Newton driver.
//
// rs_stsparse() is the GPU counterpart of call_stsparse() + dsparse()
// (src/call_stsparse.c:60-320, src/sparse.f:30-195): copy the inputs in, build
// structure / grouping / row-wise pattern once, then iterate
//   { beta = -f(y), FD Jacobian ; test 1 ; solve J x = beta ; y += x, clamp ;
//     test 2 ; new error weights }
// and hand back y with the precis / steady attributes.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <mutex>

#include "solver.cuh"

namespace rs {

// ---- error text, context ---------------------------------------------------------
static thread_local char g_err[512] = "";
std::atomic<long long> g_device_bytes{0};

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

// ---- caching device allocator (see rs_internal.cuh) ------------------------------------------
namespace {
struct Pool {
    std::mutex mu;
    std::multimap<std::pair<int, size_t>, void *> free_blocks;   // (device, class bytes) -> block
    size_t cached = 0;
    size_t max_cached = 48ull << 30;
    bool configured = false;
    void configure()
    {
        if (configured) return;
        configured = true;
        if (const char *e = getenv("RS_POOL_MAX_BYTES")) max_cached = (size_t)strtoull(e, nullptr, 10);
    }
    // size classes: 512 B steps below 1 MiB, 1 MiB steps above -- solves of the same shape ask for the same sizes
    static size_t klass(size_t bytes) { return bytes < (1u << 20) ? (bytes + 511) & ~(size_t)511 : (bytes + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1); }
    size_t drop_all_locked()
    {
        size_t freed = 0;
        for (auto &kv : free_blocks) { cudaFree(kv.second); freed += kv.first.second; }
        free_blocks.clear();
        cached = 0;
        return freed;
    }
};
Pool &pool() { static Pool *p = new Pool(); return *p; }   // never destroyed: outlives the CUDA runtime's own teardown order
}  // namespace

cudaError_t pool_alloc(void **p, size_t bytes)
{
    Pool &P = pool();
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const size_t k = Pool::klass(bytes);
    {
        std::lock_guard<std::mutex> lk(P.mu);
        P.configure();
        auto it = P.free_blocks.find({dev, k});
        if (it != P.free_blocks.end()) {
            *p = it->second;
            P.free_blocks.erase(it);
            P.cached -= k;
            return cudaSuccess;
        }
    }
    e = cudaMalloc(p, k);
    if (e == cudaErrorMemoryAllocation) {   // give the cache back and try once more
        cudaGetLastError();
        {
            std::lock_guard<std::mutex> lk(P.mu);
            P.drop_all_locked();
        }
        e = cudaMalloc(p, k);
    }
    return e;
}

void pool_free(void *p, size_t bytes)
{
    if (!p) return;
    Pool &P = pool();
    const size_t k = Pool::klass(bytes);
    // cudaFree waits for the device; a block that goes back to the cache gets the same guarantee, so whoever is
    // handed it next (possibly on another stream) cannot race with work still using it.  The cache is keyed by the
    // device that OWNS the block, not by whichever is current when it is freed.
    int dev = 0;
    cudaPointerAttributes attr;
    bool ok = cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeDevice;
    if (ok) dev = attr.device; else cudaGetLastError();
    ok = ok && cudaDeviceSynchronize() == cudaSuccess;
    std::lock_guard<std::mutex> lk(P.mu);
    P.configure();
    if (!ok || P.cached + k > P.max_cached) { cudaFree(p); return; }
    P.free_blocks.insert({{dev, k}, p});
    P.cached += k;
}

namespace {
struct PinnedPool {
    std::mutex mu;
    std::vector<void *> free_blocks;
};
PinnedPool &pinned_pool() { static PinnedPool *p = new PinnedPool(); return *p; }
}  // namespace

cudaError_t pinned_alloc(void **p)
{
    PinnedPool &P = pinned_pool();
    {
        std::lock_guard<std::mutex> lk(P.mu);
        if (!P.free_blocks.empty()) { *p = P.free_blocks.back(); P.free_blocks.pop_back(); return cudaSuccess; }
    }
    return cudaMallocHost(p, PINNED_BLOCK);
}

void pinned_free(void *p)
{
    if (!p) return;
    PinnedPool &P = pinned_pool();
    std::lock_guard<std::mutex> lk(P.mu);
    if (P.free_blocks.size() < 64) P.free_blocks.push_back(p); else cudaFreeHost(p);
}

const char *const kKernelNames[K_COUNT] = {"structure", "scan", "transpose", "jgroup", "perturb", "residual_batched",
                                           "fd_assemble", "spmv_csr", "krylov_vec", "precond_setup", "precond_apply",
                                           "newton", "halo", "misc", "spmv_grid", "mg_smooth_l0", "mg_transfer_l0",
                                           "mg_coarse_levels", "fd_rows_fused"};

void Ctx::begin(int id, double algo_bytes)
{
    bytes[id] += algo_bytes;
    Pending p;
    p.id = id;
    auto get = [&]() {
        cudaEvent_t e;
        if (!pool.empty()) { e = pool.back(); pool.pop_back(); }
        else cudaEventCreate(&e);
        return e;
    };
    p.a = get();
    p.b = get();
    cudaEventRecord(p.a, stream);
    pending.push_back(p);
}
void Ctx::end(int)
{
    cudaEventRecord(pending.back().b, stream);
    if (pending.size() >= 8192) resolve();
}
void Ctx::resolve()
{
    if (pending.empty()) return;
    cudaStreamSynchronize(stream);
    for (auto &p : pending) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) ms[p.id] += t;
        pool.push_back(p.a);
        pool.push_back(p.b);
    }
    pending.clear();
}
void Ctx::reset_stats()
{
    resolve();
    for (int i = 0; i < K_COUNT; i++) { ms[i] = 0; count[i] = 0; bytes[i] = 0; }
    launches = 0;
}
Ctx::Ctx()
{
    if (cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking) != cudaSuccess) {
        stream = nullptr;   // no device: every entry point reports RS_ENODEVICE before launching anything
        cudaGetLastError();
    }
}
Ctx::~Ctx()
{
    resolve();
    for (auto e : pool) cudaEventDestroy(e);
    if (stream) cudaStreamDestroy(stream);
}

int ensure_device()
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        set_error("no CUDA device available (%s); this library has no CPU path",
                  e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        cudaGetLastError();
        return RS_ENODEVICE;
    }
    return RS_OK;
}

// ---- Newton kernels (K9) -------------------------------------------------------------
// ewt = rtol |y| + atol with the four itol variants (errset, src/sparse.f:264-293)
__device__ __forceinline__ double ewt_of(int itol, const double *__restrict__ rtol, const double *__restrict__ atol,
                                         long long i, double yi)
{
    const double rt = (itol == 3 || itol == 4) ? rtol[i] : rtol[0];
    const double at = (itol == 2 || itol == 4) ? atol[i] : atol[0];
    return rt * fabs(yi) + at;
}

// All Newton kernels work on an index range [b, e) of global-index vectors (one range on a single GPU,
// one per species of the slab when sharded); `accumulate` chains the reductions of several ranges.
__global__ void errset_kernel(int b, int e, int itol, const double *__restrict__ rtol, const double *__restrict__ atol,
                              const double *__restrict__ y, double *__restrict__ ewt)
{
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x)
        ewt[i] = ewt_of(itol, rtol, atol, i, y[i]);
}

// test 1 (src/sparse.f:122-133): out = { sum |beta|, max |beta/ewt| }
__global__ void __launch_bounds__(256) test1_kernel(int b, int e, const double *__restrict__ beta,
                                                    const double *__restrict__ ewt, double *partials, unsigned *ticket,
                                                    double *out, int accumulate)
{
    double acc[2] = {0.0, 0.0};
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) {
        const double bi = beta[i];
        acc[0] += fabs(bi);
        acc[1] = fmax(acc[1], fabs(bi / ewt[i]));
    }
    grid_reduce<2>(acc, 2u, partials, ticket, out, accumulate != 0);
}

// update (src/sparse.f:152-163, 185): out = { max |x| } ; y += x ; clamp ; ewt(y)
__global__ void __launch_bounds__(256) update_kernel(int b, int e, int positivity, int itol, const double *__restrict__ rtol,
                                                     const double *__restrict__ atol, const double *__restrict__ x,
                                                     double *__restrict__ y, double *__restrict__ ewt, double *partials,
                                                     unsigned *ticket, double *out, int accumulate)
{
    double acc[1] = {0.0};
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) {
        const double xi = x[i];
        acc[0] = fmax(acc[0], fabs(xi));
        double yn = y[i] + xi;
        if (positivity) yn = fmax(0.0, yn);
        y[i] = yn;
        ewt[i] = ewt_of(itol, rtol, atol, i, yn);
    }
    grid_reduce<1>(acc, 1u, partials, ticket, out, accumulate != 0);
}

// index-vector form of `positive` (src/sparse.f:159-163); only indices inside [b, e) are touched
__global__ void clamp_indices_kernel(int b, int e, int npos, const int *__restrict__ pos, int itol,
                                     const double *__restrict__ rtol, const double *__restrict__ atol,
                                     double *__restrict__ y, double *__restrict__ ewt)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= npos) return;
    const int i = pos[q] - 1;
    if (i < b || i >= e) return;
    const double yn = fmax(0.0, y[i]);
    y[i] = yn;
    ewt[i] = ewt_of(itol, rtol, atol, i, yn);
}

// test 2's extra evaluation (src/sparse.f:165-183): out = { sum |f| }
__global__ void __launch_bounds__(256) sumabs_kernel(int b, int e, const double *__restrict__ f, double *partials,
                                                     unsigned *ticket, double *out, int accumulate)
{
    double acc[1] = {0.0};
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x)
        acc[0] += fabs(f[i]);
    grid_reduce<1>(acc, 0u, partials, ticket, out, accumulate != 0);
}

}  // namespace rs

using namespace rs;

// a pair of CUDA events that does not leak on an early return
struct EventPair {
    cudaEvent_t a = nullptr, b = nullptr;
    int create() { RS_CUDA(cudaEventCreate(&a)); RS_CUDA(cudaEventCreate(&b)); return RS_OK; }
    ~EventPair() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
};

// ---- the handle ----------------------------------------------------------------------------
struct rs_solver {
    Ctx ctx;
    int n = 0, nnz = 0, ngp = 0, maxg = 0, type = 0, imethod = 1;
    GridDesc grid;
    const rs_model_desc *model = nullptr;
    double time = 0.0, ctol = 1e-8;
    int itol = 1, maxiter = 100, Pos = 0, nout = 0, npos = 0, verbose = 0;
    double min_atol = 0.0;
    rs_krylov_control ctl = {RS_PRECOND_AUTO, 0, 0.0, 0.0, 0};
    DevBuf<int> ian, jan, rowptr, col, group_of, pos;
    DevBuf<double> val, y, beta, ewt, x, atol, rtol, ftmp;
    DevBuf<double> partials, red;
    DevBuf<unsigned> ticket;
    double *h_red = nullptr;
    FdWork fd;
    Krylov kry;
    Ranges own, touch;          // rows this rank solves for / unknowns it reads (own + halo planes)
    HaloPlan halo;
    int slab_j0 = 0, slab_j1 = -1, plane = 0;   // slab of the slowest grid index (j1 < 0: not sharded)
    bool structure_only = false;
    std::vector<double> precis;
    int niter = 0, steady = 0, krylov_iters = 0;
    double t_setup = 0.0, t_newton = 0.0;
    long long setup_launches = 0;
    ~rs_solver() { pinned_free(h_red); }
};

// parameters reach the model through its initialiser, as in src/call_stsparse.c:231-240
static thread_local const double *tl_parms = nullptr;
static thread_local int tl_nparms = 0, tl_parm_err = 0;
static void steadyparms_cb(int *N, double *dst)
{
    if (*N != tl_nparms) {  // src/steady_utils.c:48-53
        set_error("Confusion over the length of parms: passed to solver %d; in model %d", tl_nparms, *N);
        tl_parm_err = 1;
        return;
    }
    for (int i = 0; i < *N; i++) dst[i] = tl_parms[i];
}

// ... and the forcing function values through initforc, src/call_stsparse.c:242-251, src/steady_utils.c:62-77
static void steadyforcs_cb(int *N, double *dst)
{
    if (*N != tl_nparms) {
        set_error("Confusion over the length of forcs: passed to solver %d; in model %d", tl_nparms, *N);
        tl_parm_err = 1;
        return;
    }
    for (int i = 0; i < *N; i++) dst[i] = tl_parms[i];
}

static int init_model(const rs_model_desc *m, const double *initpar, int ninitpar, const double *forcs = nullptr,
                      int nforcs = 0)
{
    if (!m || m->abi != RS_MODEL_ABI || !m->launch) {
        set_error("`func' must be a compiled device model (rs_model_desc with ABI %d)", RS_MODEL_ABI);
        return RS_EMODEL;
    }
    if (m->initfunc && initpar) {
        tl_parms = initpar; tl_nparms = ninitpar; tl_parm_err = 0;
        m->initfunc(steadyparms_cb);
        if (tl_parm_err) return RS_EMODEL;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("model initialiser failed: %s", cudaGetErrorString(e)); return RS_ENODEVICE; }
    }
    if (forcs && nforcs > 0) {
        if (!m->initforc) {
            set_error("forcings were given but the model exports no forcing initialiser (initforc)");
            return RS_EMODEL;
        }
        tl_parms = forcs; tl_nparms = nforcs; tl_parm_err = 0;
        m->initforc(steadyforcs_cb);
        if (tl_parm_err) return RS_EMODEL;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("forcing initialiser failed: %s", cudaGetErrorString(e)); return RS_ENODEVICE; }
    }
    return RS_OK;
}

static int grid_from_nnz(int type, const int *nnzv, int nnnz, int n, GridDesc &g)
{
    int at = std::abs(type);
    const bool mapped = at == RS_TYPE_2DMAP || at == RS_TYPE_3DMAP;
    if (mapped) at = at == RS_TYPE_2DMAP ? RS_TYPE_2D : RS_TYPE_3D;
    g = GridDesc();
    g.type = at;
    const int need = at == RS_TYPE_1D ? 4 : at == RS_TYPE_2D ? 6 : 8;
    if (nnnz < need + (mapped ? 1 : 0) || (mapped && (nnzv[need] < 0 || nnnz < need + 1 + nnzv[need]))) {
        set_error("nnz vector too short for sparsity type %d", type);
        return RS_EINVAL;
    }
    g.nspec = nnzv[1];
    if (at == RS_TYPE_1D) { g.d[0] = nnzv[2]; g.cyc[0] = nnzv[3]; g.ncell = g.d[0]; }
    if (at == RS_TYPE_2D) { g.d[0] = nnzv[2]; g.d[1] = nnzv[3]; g.cyc[0] = nnzv[4]; g.cyc[1] = nnzv[5]; g.ncell = g.d[0] * g.d[1]; }
    if (at == RS_TYPE_3D) {
        g.d[0] = nnzv[2]; g.d[1] = nnzv[3]; g.d[2] = nnzv[4];
        g.cyc[0] = nnzv[5]; g.cyc[1] = nnzv[6]; g.cyc[2] = nnzv[7];
        g.ncell = g.d[0] * g.d[1] * g.d[2];
    }
    if (mapped) return RS_OK;   // build_mapped_structure checks the map against the grid
    if (g.nspec <= 0 || g.ncell <= 0 || (long long)g.nspec * g.ncell != n) {
        set_error("cannot generate sparse jacobian - N and nspec not compatible");
        return RS_ESTRUCT;
    }
    return RS_OK;
}

static int check_args(const rs_stsparse_args *a)
{
    if (!a || !a->y || a->neq <= 0) { set_error("`y' must be numeric and non-empty"); return RS_EINVAL; }
    if (a->maxiter < 1) { set_error("maxiter must be >=1"); return RS_EINVAL; }
    if (!a->atol || !a->rtol || a->natol < 1 || a->nrtol < 1) { set_error("`atol' and `rtol' must be numeric"); return RS_EINVAL; }
    if (a->natol > 1 && a->natol != a->neq) { set_error("`atol' must either be a scalar, or as long as `y'"); return RS_EINVAL; }
    if (a->nrtol > 1 && a->nrtol != a->neq) { set_error("`rtol' must either be a scalar, or as long as `y'"); return RS_EINVAL; }
    // 1 yale, 2 ilut, 3 ilutp, 10 yale + return the structure (R/stodes.R:314-329): all served by the Krylov solve
    if (a->imethod != 1 && a->imethod != 2 && a->imethod != 3 && a->imethod != 10) {
        set_error("`spmethod' must be one of `yale', `ilut' or `ilutp' (imethod 1, 2, 3 or 10)");
        return RS_EINVAL;
    }
    const int at = std::abs(a->Type);
    if (at != RS_TYPE_USER && at != RS_TYPE_DETECT && at != RS_TYPE_1D && at != RS_TYPE_2D && at != RS_TYPE_3D &&
        at != RS_TYPE_2DMAP && at != RS_TYPE_3DMAP) {
        set_error("cannot run stodes: sparsetype not known (Type %d is outside this path)", a->Type);
        return RS_EINVAL;
    }
    if (!a->nnz || a->nnnz < 1) { set_error("nnz must be given"); return RS_EINVAL; }
    if (a->Type == 0 && (!a->ian || !a->jan)) { set_error("'inz' must be specified if 'sparsetype' = 'sparseusr'"); return RS_EINVAL; }
    for (int q = 0; q < a->npositive && a->Pos == 0 && a->npositive > 1; q++)
        if (a->positive[q] < 1 || a->positive[q] > a->neq) {
            set_error("the elements of 'positive' should be >0 and < the number of state variables");
            return RS_EINVAL;
        }
    return RS_OK;
}

extern "C" int rs_solver_create(const rs_stsparse_args *a, rs_solver **out)
{
    if (!out) return RS_EINVAL;
    *out = nullptr;
    RS_TRY(check_args(a));
    RS_TRY(ensure_device());
    std::unique_ptr<rs_solver> S(new rs_solver());
    rs_solver &s = *S;
    Ctx &c = s.ctx;
    s.n = a->neq; s.type = a->Type; s.imethod = a->imethod; s.time = a->time; s.ctol = a->ctol;
    s.itol = a->itol; s.maxiter = a->maxiter; s.Pos = a->Pos; s.nout = a->nout; s.verbose = a->verbose;
    s.maxg = a->ngp > 0 ? a->ngp : a->neq + 1;
    s.model = a->func;
    s.structure_only = a->Type < 0;
    if (a->control) s.ctl = *a->control;
    const int n = s.n;
    EventPair ev;
    RS_TRY(ev.create());
    cudaEvent_t e0 = ev.a, e1 = ev.b;
    RS_CUDA(cudaEventRecord(e0, c.stream));

    // RS_DEBUG_SETUP=1: wall-clock per set-up phase on stderr (each mark drains the stream)
    static const bool debug_setup = getenv("RS_DEBUG_SETUP") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto mark = [&](const char *what) {
        if (!debug_setup) return;
        cudaStreamSynchronize(c.stream);
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[setup] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    RS_TRY(s.y.alloc(n));
    // Grid types: structure and grouping do not read y, so a large state could cross PCIe on a stream of its own beside
    // them; the solver's stream then waits for it before the first kernel that reads y.  Opt-in (RS_SIDE_COPY=1): on
    // the B200 boxes this was measured on the copy did not overlap the set-up kernels at all -- 11.8 ms of set-up at
    // 256^3 with it against 10.95 ms with the upload simply first on the solver's stream (2.4 + 8.5 ms).  The guard
    // drains the copy on every return path: y must not go back to the pool under it.
    struct SideCopy {
        cudaStream_t st = nullptr; cudaEvent_t done = nullptr; bool pending = false;
        ~SideCopy() { if (pending) cudaStreamSynchronize(st); if (done) cudaEventDestroy(done); if (st) cudaStreamDestroy(st); }
    } side;
    const bool grid_type = std::abs(a->Type) == RS_TYPE_1D || std::abs(a->Type) == RS_TYPE_2D || std::abs(a->Type) == RS_TYPE_3D;
    static const bool side_copy = getenv("RS_SIDE_COPY") != nullptr;
    if (side_copy && grid_type && n >= (1 << 18) && !debug_setup) {
        RS_CUDA(cudaStreamCreateWithFlags(&side.st, cudaStreamNonBlocking));
        RS_CUDA(cudaEventCreateWithFlags(&side.done, cudaEventDisableTiming));
        side.pending = true;
        RS_TRY(s.y.upload(a->y, n, side.st));
        RS_CUDA(cudaEventRecord(side.done, side.st));
    } else
        RS_TRY(s.y.upload(a->y, n, c.stream));
    mark("state upload");
    // structure (xsparsestruct, src/sparse.f:300-455)
    if (std::abs(a->Type) == RS_TYPE_DETECT) {
        RS_TRY(init_model(a->func, a->initpar, a->ninitpar, a->forcs, a->nforcs));
        RS_TRY(detect_structure(c, a->func, n, a->time, s.y.p, a->nout, a->rpar, a->nrpar, a->ipar, a->nipar, a->nnz[0], s.ian,
                                s.jan, &s.nnz));
    } else if (std::abs(a->Type) == RS_TYPE_USER) {
        s.nnz = a->nnz[0];
        RS_TRY(s.ian.alloc((size_t)n + 1));
        RS_TRY(s.jan.alloc((size_t)s.nnz));
        RS_TRY(s.ian.upload(a->ian, (size_t)n + 1, c.stream));
        RS_TRY(s.jan.upload(a->jan, (size_t)s.nnz, c.stream));
    } else if (std::abs(a->Type) == RS_TYPE_2DMAP || std::abs(a->Type) == RS_TYPE_3DMAP) {
        GridDesc full;
        RS_TRY(grid_from_nnz(a->Type, a->nnz, a->nnnz, n, full));
        const int at = std::abs(a->Type) == RS_TYPE_2DMAP ? 6 : 8;      // TotN, then the map (src/call_stsparse.c:196-206)
        RS_TRY(build_mapped_structure(c, full, n, a->nnz[at], a->nnz + at + 1, s.ian, s.jan, &s.nnz));
        if (s.nnz >= a->nnz[0]) {   // updatejan stops once the write position passes nnz, src/testMap.f:17-20
            set_error("cannot generate sparse jacobian - not enough room for nonzeros");
            return RS_ESTRUCT;
        }
        // s.grid stays "no grid knowledge": grouping, preconditioner and sharding take the general-structure routes
    } else {
        RS_TRY(grid_from_nnz(a->Type, a->nnz, a->nnnz, n, s.grid));
        RS_TRY(build_grid_structure(c, s.grid, n, s.ian, s.jan, &s.nnz));
        if (s.nnz > a->nnz[0] && std::abs(a->Type) == RS_TYPE_3D) {  // interact's capacity check, src/sparsecoeff.f:193-196
            set_error("cannot generate sparse jacobian - not enough room for nonzeros");
            return RS_ESTRUCT;
        }
    }
    mark("structure (ian, jan)");
    if (!s.structure_only) {
        RS_TRY(init_model(a->func, a->initpar, a->ninitpar, a->forcs, a->nforcs));
        if (grid_pattern_applicable(s.grid)) RS_TRY(build_csr_pattern_grid(c, s.grid, n, s.nnz, s.ian.p, s.jan.p, s.rowptr, s.col));
        else RS_TRY(build_csr_pattern(c, n, s.nnz, s.ian.p, s.jan.p, s.rowptr, s.col));
        mark("row-wise pattern");
        RS_TRY(jgroup_device(c, s.grid, n, s.nnz, s.ian.p, s.jan.p, s.rowptr.p, s.col.p, s.maxg, s.group_of, &s.ngp));
        mark("column grouping");
        RS_TRY(s.val.alloc((size_t)s.nnz));
        RS_TRY(s.beta.alloc(n)); RS_TRY(s.ewt.alloc(n)); RS_TRY(s.x.alloc(n)); RS_TRY(s.ftmp.alloc(n));
        RS_TRY(s.atol.alloc(a->natol)); RS_TRY(s.rtol.alloc(a->nrtol));
        RS_TRY(s.atol.upload(a->atol, a->natol, c.stream));
        RS_TRY(s.rtol.upload(a->rtol, a->nrtol, c.stream));
        s.min_atol = *std::min_element(a->atol, a->atol + a->natol);
        if (a->Pos == 0 && a->npositive > 1) {
            s.npos = a->npositive;
            RS_TRY(s.pos.alloc(s.npos));
            RS_TRY(s.pos.upload(a->positive, s.npos, c.stream));
        }
        RS_TRY(s.partials.alloc(4 * 148 * 16));
        RS_TRY(s.red.alloc(8));
        RS_TRY(s.ticket.alloc(4));
        RS_TRY(s.ticket.zero(c.stream));
        RS_CUDA(pinned_alloc((void **)&s.h_red));
        // who owns what: one range [0, n) on a single GPU; planes [j0, j1) of every species when sharded
        s.own.n = 1; s.own.b[0] = 0; s.own.e[0] = n;
        s.touch = s.own;
        if (comm_active()) {
            const GridDesc &g = s.grid;
            const int nr = comm_nranks(), rk = comm_rank();
            if (!g.type) { set_error("a sharded solve needs a grid sparsity type (1D/2D/3D)"); return RS_EINVAL; }
            if (g.cyc[0]) { set_error("a cyclic boundary in the sharded (slowest) dimension is not supported"); return RS_EINVAL; }
            if (nr > g.d[0]) { set_error("more ranks (%d) than planes (%d) in the sharded dimension", nr, g.d[0]); return RS_EINVAL; }
            if (g.nspec > MAXR) { set_error("more than %d species in a sharded solve", MAXR); return RS_ETOOBIG; }
            s.plane = g.ncell / g.d[0];
            rs_slab_bounds(g.d[0], nr, rk, &s.slab_j0, &s.slab_j1);
            const int jl = std::max(s.slab_j0 - 1, 0), jh = std::min(s.slab_j1 + 1, g.d[0]);
            s.own.n = s.touch.n = g.nspec;
            for (int q = 0; q < g.nspec; q++) {
                s.own.b[q] = q * g.ncell + s.slab_j0 * s.plane;   s.own.e[q] = q * g.ncell + s.slab_j1 * s.plane;
                s.touch.b[q] = q * g.ncell + jl * s.plane;        s.touch.e[q] = q * g.ncell + jh * s.plane;
            }
            s.halo.nspec = g.nspec; s.halo.ncell = g.ncell; s.halo.plane = s.plane;
            s.halo.j0 = s.slab_j0; s.halo.j1 = s.slab_j1;
            s.halo.has_lo = rk > 0; s.halo.has_hi = rk < nr - 1;
        }
        mark("newton vectors");
        if (side.pending) RS_CUDA(cudaStreamWaitEvent(c.stream, side.done, 0));
        RS_TRY(fd_setup(c, s.fd, n, s.ngp, a->nout, a->rpar, a->nrpar, a->ipar, a->nipar, s.slab_j0, s.slab_j1));
        mark("fd work arrays");
        RS_TRY(krylov_setup(c, s.kry, n, s.nnz, s.rowptr.p, s.col.p, s.val.p, s.ctl.precond, s.grid.type ? s.grid.nspec : 1,
                            s.grid.type ? s.grid.ncell : n, &s.ctl, &s.own, comm_active() ? &s.halo : nullptr,
                            s.grid.type ? &s.grid : nullptr));
        mark("krylov set-up");
        // the FD assembly shares the SpMV's knowledge of the grid (pattern verified row by row in krylov_setup)
        s.fd.gridcols = s.kry.gridcols; s.fd.gd1 = s.kry.gd1; s.fd.gd2 = s.kry.gd2; s.fd.gd3 = s.kry.gd3; s.fd.gncell = s.kry.ncell;
    }
    RS_CUDA(cudaEventRecord(e1, c.stream));
    RS_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    s.t_setup = ms * 1e-3;
    s.setup_launches = c.launches;
    *out = S.release();
    return RS_OK;
}

extern "C" int rs_solver_set_state(rs_solver *s, const double *y, int is_device)
{
    if (!s || !y) return RS_EINVAL;
    RS_CUDA(cudaMemcpyAsync(s->y.p, y, sizeof(double) * (size_t)s->n, is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                            s->ctx.stream));
    return RS_OK;
}

extern "C" double *rs_solver_state_device(rs_solver *s) { return s ? s->y.p : nullptr; }

extern "C" int rs_solver_get_state(rs_solver *s, double *y_host)
{
    if (!s || !y_host) return RS_EINVAL;
    if (comm_active() && !s->structure_only)
        RS_TRY(comm_gather_slabs(s->ctx, s->y.p, s->grid.nspec, s->grid.ncell, s->plane, s->grid.d[0]));
    RS_TRY(s->y.download(y_host, s->n, s->ctx.stream));
    RS_CUDA(cudaStreamSynchronize(s->ctx.stream));
    return RS_OK;
}

static int read_red(rs_solver &s, int count)
{
    RS_CUDA(cudaMemcpyAsync(s.h_red, s.red.p, sizeof(double) * count, cudaMemcpyDeviceToHost, s.ctx.stream));
    RS_CUDA(cudaStreamSynchronize(s.ctx.stream));
    return RS_OK;
}

// The Newton loop, src/sparse.f:114-188.
extern "C" int rs_solver_run(rs_solver *sp, rs_stsparse_result *meta)
{
    if (!sp) return RS_EINVAL;
    rs_solver &s = *sp;
    Ctx &c = s.ctx;
    if (s.structure_only) { s.niter = 1; s.steady = 0; s.precis.assign(1, 0.0); return RS_OK; }
    const int n = s.n;
    const Ranges &own = s.own;
    const double rows = (double)own.total();
    cudaStream_t st = c.stream;
    EventPair ev;
    RS_TRY(ev.create());
    cudaEvent_t e0 = ev.a, e1 = ev.b;
    RS_CUDA(cudaEventRecord(e0, st));
    s.precis.assign(s.maxiter, 0.0);
    s.niter = 0; s.steady = 0; s.krylov_iters = 0;
    const long long launches0 = c.launches;
    {
        Launch L(c, K_NEWTON, 16.0 * rows);
        for (int q = 0; q < own.n; q++)
            errset_kernel<<<grid_for(own.e[q] - own.b[q], 256), 256, 0, st>>>(own.b[q], own.e[q], s.itol, s.rtol.p, s.atol.p, s.y.p, s.ewt.p);
    }
    for (int it = 1; it <= s.maxiter; it++) {
        s.niter = it;
        RS_TRY(comm_halo_exchange(c, s.y.p, s.halo));   // the residual reads one plane beyond the slab
        // xSparseJacob's first model call (beta = -f(y)); its ngp perturbed calls follow below, after test 1,
        // and only if a Newton step is taken: the reference computes that last Jacobian and drops it
        RS_TRY(fd_base(c, s.model, s.fd, n, s.time, s.y.p, s.beta.p, own));
        {
            Launch L(c, K_NEWTON, 16.0 * rows);
            for (int q = 0; q < own.n; q++)
                test1_kernel<<<grid_for(own.e[q] - own.b[q], 256), 256, 0, st>>>(own.b[q], own.e[q], s.beta.p, s.ewt.p, s.partials.p,
                                                                                 s.ticket.p, s.red.p, q > 0);
        }
        RS_TRY(comm_allreduce_sum(c, s.red.p, 1));
        RS_TRY(comm_allreduce_max(c, s.red.p + 1, 1));
        RS_TRY(read_red(s, 2));
        s.precis[it - 1] = s.h_red[0] / n;
        if (!(s.h_red[0] == s.h_red[0])) { set_error("model returned NaN in the rate of change"); return RS_ESINGULAR; }
        if (s.h_red[1] <= 1.0) { s.steady = 1; break; }

        RS_TRY(fd_groups(c, s.model, s.fd, n, s.nnz, s.ngp, s.time, s.y.p, s.group_of.p, s.rowptr.p, s.col.p, s.val.p,
                         s.beta.p, own, s.touch));
        RS_TRY(krylov_precond_update(c, s.kry));
        int kit = 0;
        double relres = 0.0;
        // Absolute floor of the linear residual: once its RMS is a small fraction of the smallest error
        // weight, the next residual test (max |f/ewt| <= 1) passes and further Krylov work is wasted.
        bool lin_converged = false;
        RS_TRY(krylov_solve(c, s.kry, s.beta.p, s.x.p, s.kry.lin_ewt_factor * s.min_atol * std::sqrt((double)n), &kit,
                            &relres, &lin_converged));
        s.krylov_iters += kit;
        // An inexact Newton step is fine; no step at all is not: a solve that barely reduced the residual
        // would otherwise look like convergence to test 2 (max|x| <= ctol).  The reference's direct solve
        // cannot fail this way -- it stops with "zero pivot" instead (src/sparse.f:2376).
        // A solve that stopped on its own target is a step, whatever its relative residual: close to the absolute
        // floor ||beta|| itself is within a small factor of the target.
        if (!lin_converged && !(relres <= 0.1)) {
            set_error("sparse solver: the Krylov solve of Newton iteration %d stalled at relative residual %.3e after %d "
                      "iterations (matrix too ill-conditioned or indefinite for the chosen preconditioner)", it, relres, kit);
            return RS_ESINGULAR;
        }
        if (s.verbose && comm_rank() == 0)
            fprintf(stderr, "[rootsolve_b200] newton %d: mean|f|=%.6e  krylov its=%d relres=%.3e\n", it, s.precis[it - 1], kit, relres);
        {
            Launch L(c, K_NEWTON, 48.0 * rows);
            for (int q = 0; q < own.n; q++)
                update_kernel<<<grid_for(own.e[q] - own.b[q], 256), 256, 0, st>>>(own.b[q], own.e[q], s.Pos > 0 ? 1 : 0, s.itol, s.rtol.p,
                                                                                  s.atol.p, s.x.p, s.y.p, s.ewt.p, s.partials.p,
                                                                                  s.ticket.p, s.red.p, q > 0);
        }
        if (s.Pos <= 0 && s.npos > 1) {
            Launch L(c, K_NEWTON, 28.0 * s.npos);
            for (int q = 0; q < own.n; q++)
                clamp_indices_kernel<<<(s.npos + 255) / 256, 256, 0, st>>>(own.b[q], own.e[q], s.npos, s.pos.p, s.itol, s.rtol.p,
                                                                           s.atol.p, s.y.p, s.ewt.p);
        }
        RS_TRY(comm_allreduce_max(c, s.red.p, 1));
        RS_TRY(read_red(s, 1));
        const double relchange = s.h_red[0];
        if (relchange <= s.ctol) {
            if (it < s.maxiter) {
                RS_TRY(comm_halo_exchange(c, s.y.p, s.halo));
                RS_TRY(residual_only(c, s.model, s.fd, n, s.time, s.y.p, s.ftmp.p, own.total()));
                {
                    Launch L(c, K_NEWTON, 8.0 * rows);
                    for (int q = 0; q < own.n; q++)
                        sumabs_kernel<<<grid_for(own.e[q] - own.b[q], 256), 256, 0, st>>>(own.b[q], own.e[q], s.ftmp.p, s.partials.p,
                                                                                          s.ticket.p, s.red.p, q > 0);
                }
                RS_TRY(comm_allreduce_sum(c, s.red.p, 1));
                RS_TRY(read_red(s, 1));
                s.precis[it] = s.h_red[0];  // a sum, not a mean: src/sparse.f:177
                s.niter = it + 1;
            }
            s.steady = 1;
            break;
        }
    }
    RS_CUDA(cudaGetLastError());
    RS_CUDA(cudaEventRecord(e1, st));
    RS_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    s.t_newton = ms * 1e-3;
    if (meta) {
        meta->niter = s.niter;
        meta->steady[0] = s.steady; meta->steady[1] = s.nnz; meta->steady[2] = s.ngp;
        meta->steady[3] = (int)std::min<long long>(g_device_bytes.load() / 8, 2147483647LL);
        meta->krylov_iters = s.krylov_iters;
        meta->kernel_launches = (int)(c.launches - launches0);
        meta->t_setup_s = s.t_setup;
        meta->t_newton_s = s.t_newton;
        meta->spmv_kernel = s.kry.gridcols ? 2 : s.kry.tma_ok ? 1 : 0;
        meta->precond = s.kry.mg_on ? RS_PRECOND_MG : s.kry.btd_on ? RS_PRECOND_BTD : s.kry.precond;
        meta->krylov_method = s.kry.use_cg ? 1 : 0;
        if (meta->precis) for (int q = 0; q < s.niter && q < s.maxiter; q++) meta->precis[q] = s.precis[q];
    }
    return RS_OK;
}

extern "C" int rs_solver_set_profiling(rs_solver *s, int on)
{
    if (!s) return RS_EINVAL;
    s->ctx.resolve();
    s->ctx.profile = on != 0;
    s->ctx.reset_stats();
    return RS_OK;
}

extern "C" int rs_solver_kernel_stats(rs_solver *s, int cap, const char **names, double *ms, long long *launches, double *bytes)
{
    if (!s) return 0;
    s->ctx.resolve();
    int k = 0;
    for (int i = 0; i < K_COUNT && k < cap; i++) {
        if (s->ctx.count[i] == 0) continue;
        names[k] = kKernelNames[i];
        ms[k] = s->ctx.ms[i];
        launches[k] = s->ctx.count[i];
        bytes[k] = s->ctx.bytes[i] / (double)s->ctx.count[i];   // average algorithmic bytes per launch of the class
        k++;
    }
    return k;
}

extern "C" void rs_solver_destroy(rs_solver *s) { delete s; }

// Fills the caller's result buffers from a finished handle (src/call_stsparse.c:278-317).
static int pack_result(rs_solver &s, const rs_stsparse_args *a, rs_stsparse_result *r)
{
    Ctx &c = s.ctx;
    const int n = s.n;
    if (r->y) {
        if (comm_active() && !s.structure_only)   // every rank returns the whole state, as the R call does
            RS_TRY(comm_gather_slabs(c, s.y.p, s.grid.nspec, s.grid.ncell, s.plane, s.grid.d[0]));
        RS_TRY(s.y.download(r->y, n, c.stream));
        if (a->nout > 0 && !s.structure_only) {  // one more derivs call for the outputs, src/call_stsparse.c:281-286
            RS_TRY(residual_only(c, s.model, s.fd, n, s.time, s.y.p, s.ftmp.p, s.own.total()));
            RS_CUDA(cudaMemcpyAsync(r->y + n, s.fd.yout.p, sizeof(double) * a->nout, cudaMemcpyDeviceToHost, c.stream));
        }
    }
    if (r->ian) RS_TRY(s.ian.download(r->ian, (size_t)n + 1, c.stream));
    if (r->jan) RS_TRY(s.jan.download(r->jan, (size_t)std::min(s.nnz, a->nnz[0]), c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    if (!s.structure_only && (r->igp || r->jgp)) {
        if (r->igp) for (int q = 0; q <= s.maxg; q++) r->igp[q] = 0;
        RS_TRY(jgroup_lists(c, n, s.ngp, s.group_of.p, r->igp, r->jgp));
    }
    r->niter = s.niter;
    r->steady[0] = s.structure_only ? 0 : s.steady;
    r->steady[1] = s.nnz;
    r->steady[2] = s.ngp;
    if (r->precis) for (int q = 0; q < s.niter && q < a->maxiter; q++) r->precis[q] = s.precis[q];
    return RS_OK;
}

extern "C" int rs_stsparse(const rs_stsparse_args *a, rs_stsparse_result *r)
{
    if (!r) { set_error("result must not be NULL"); return RS_EINVAL; }
    rs_solver *s = nullptr;
    int rc = rs_solver_create(a, &s);
    if (rc != RS_OK) return rc;
    rc = rs_solver_run(s, r);
    if (rc == RS_OK) rc = pack_result(*s, a, r);
    if (rc == RS_OK) {
        r->kernel_launches = (int)s->ctx.launches;
        r->t_setup_s = s->t_setup;
        r->t_newton_s = s->t_newton;
        r->krylov_iters = s->krylov_iters;
        r->steady[3] = (int)std::min<long long>(g_device_bytes.load() / 8, 2147483647LL);
    }
    rs_solver_destroy(s);
    return rc;
}

// ---- building blocks for the parity tests ---------------------------------------------------------
extern "C" int rs_sparse_structure(int type, int neq, int nspec, const int *dims, const int *cyclic, int *nnz, int *ian, int *jan)
{
    if (!dims || !nnz || !ian || !jan) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    Ctx c;
    GridDesc g;
    g.type = type; g.nspec = nspec;
    const int nd = type == RS_TYPE_1D ? 1 : type == RS_TYPE_2D ? 2 : type == RS_TYPE_3D ? 3 : 0;
    if (!nd) { set_error("type must be RS_TYPE_1D, 2D or 3D"); return RS_EINVAL; }
    long long cells = 1;
    for (int q = 0; q < nd; q++) { g.d[q] = dims[q]; g.cyc[q] = cyclic ? cyclic[q] : 0; cells *= dims[q]; }
    g.ncell = (int)cells;
    DevBuf<int> d_ian, d_jan;
    int z = 0;
    RS_TRY(build_grid_structure(c, g, neq, d_ian, d_jan, &z));
    if (z > *nnz) { set_error("cannot generate sparse jacobian - not enough room for nonzeros"); return RS_ESTRUCT; }
    RS_TRY(d_ian.download(ian, (size_t)neq + 1, c.stream));
    RS_TRY(d_jan.download(jan, (size_t)z, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    *nnz = z;
    return RS_OK;
}

extern "C" int rs_sparse_structure_map(int type, int neq, int nspec, const int *dims, const int *cyclic, const int *pres,
                                       int totn, int *nnz, int *ian, int *jan)
{
    if (!dims || !pres || !nnz || !ian || !jan) { set_error("null argument"); return RS_EINVAL; }
    if (type != RS_TYPE_2DMAP && type != RS_TYPE_3DMAP) { set_error("type must be RS_TYPE_2DMAP or RS_TYPE_3DMAP"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    Ctx c;
    GridDesc g;
    g.type = type == RS_TYPE_2DMAP ? RS_TYPE_2D : RS_TYPE_3D; g.nspec = nspec;
    const int nd = type == RS_TYPE_2DMAP ? 2 : 3;
    long long cells = 1;
    for (int q = 0; q < nd; q++) { g.d[q] = dims[q]; g.cyc[q] = cyclic ? cyclic[q] : 0; cells *= dims[q]; }
    g.ncell = (int)cells;
    DevBuf<int> d_ian, d_jan;
    int z = 0;
    RS_TRY(build_mapped_structure(c, g, neq, totn, pres, d_ian, d_jan, &z));
    if (z >= *nnz) { set_error("cannot generate sparse jacobian - not enough room for nonzeros"); return RS_ESTRUCT; }
    RS_TRY(d_ian.download(ian, (size_t)neq + 1, c.stream));
    RS_TRY(d_jan.download(jan, (size_t)z, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    *nnz = z;
    return RS_OK;
}

extern "C" int rs_jgroup(int neq, const int *ian, const int *jan, int maxg, const int *grid_hint, int *ngp, int *igp, int *jgp)
{
    if (!ian || !jan || !ngp) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    Ctx c;
    const int nnz = ian[neq] - 1;
    DevBuf<int> d_ian, d_jan, rowptr, col, group_of;
    RS_TRY(d_ian.alloc((size_t)neq + 1)); RS_TRY(d_jan.alloc((size_t)nnz));
    RS_TRY(d_ian.upload(ian, (size_t)neq + 1, c.stream));
    RS_TRY(d_jan.upload(jan, (size_t)nnz, c.stream));
    GridDesc g;
    if (grid_hint && grid_hint[0]) {
        g.type = grid_hint[0]; g.nspec = grid_hint[1];
        for (int q = 0; q < 3; q++) { g.d[q] = grid_hint[2 + q] > 0 ? grid_hint[2 + q] : 1; g.cyc[q] = grid_hint[5 + q]; }
        g.ncell = g.d[0] * g.d[1] * g.d[2];
    }
    // the rounds variant walks rows -> needs the row-wise pattern; duplicates are legal for grouping
    int rc = build_csr_pattern(c, neq, nnz, d_ian.p, d_jan.p, rowptr, col);
    if (rc != RS_OK && rc != RS_EDUPLICATE) return rc;
    RS_TRY(jgroup_device(c, g, neq, nnz, d_ian.p, d_jan.p, rowptr.p, col.p, maxg, group_of, ngp));
    if (igp) for (int q = 0; q <= maxg; q++) igp[q] = 0;
    RS_TRY(jgroup_lists(c, neq, *ngp, group_of.p, igp, jgp));
    return RS_OK;
}

static int group_of_from_lists(int neq, int ngp, const int *igp, const int *jgp, std::vector<int> &g)
{
    g.assign((size_t)neq, -1);
    for (int q = 0; q < ngp; q++) {
        // igp(ngp+1) is left at 0 by the reference when ngp == neq (see jgroup_lists)
        const int stop = (q == ngp - 1 && igp[q + 1] == 0) ? neq + 1 : igp[q + 1];
        for (int k = igp[q]; k < stop; k++) {
            const int j = jgp[k - 1];
            if (j < 1 || j > neq) { set_error("jgp out of range"); return RS_EINVAL; }
            g[j - 1] = q;
        }
    }
    for (int j = 0; j < neq; j++)
        if (g[j] < 0) { set_error("column %d is in no group", j + 1); return RS_EINVAL; }
    return RS_OK;
}

extern "C" int rs_fd_jacobian(const rs_model_desc *func, const double *initpar, int ninitpar, int neq, double time,
                              const double *y, const int *ian, const int *jan, const int *igp, const int *jgp, int ngp,
                              const double *rpar, int nrpar, const int *ipar, int nipar, double *beta, double *a)
{
    if (!y || !ian || !jan || !igp || !jgp) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    RS_TRY(init_model(func, initpar, ninitpar));
    Ctx c;
    const int nnz = ian[neq] - 1;
    std::vector<int> g;
    RS_TRY(group_of_from_lists(neq, ngp, igp, jgp, g));
    DevBuf<int> d_ian, d_jan, rowptr, col, group_of;
    DevBuf<double> d_y, d_val, d_beta, d_a;
    RS_TRY(d_ian.alloc((size_t)neq + 1)); RS_TRY(d_jan.alloc((size_t)nnz)); RS_TRY(group_of.alloc(neq));
    RS_TRY(d_y.alloc(neq)); RS_TRY(d_val.alloc((size_t)nnz)); RS_TRY(d_beta.alloc(neq)); RS_TRY(d_a.alloc((size_t)nnz));
    RS_TRY(d_ian.upload(ian, (size_t)neq + 1, c.stream));
    RS_TRY(d_jan.upload(jan, (size_t)nnz, c.stream));
    RS_TRY(group_of.upload(g.data(), neq, c.stream));
    RS_TRY(d_y.upload(y, neq, c.stream));
    RS_TRY(build_csr_pattern(c, neq, nnz, d_ian.p, d_jan.p, rowptr, col));
    FdWork w;
    RS_TRY(fd_setup(c, w, neq, ngp, 0, rpar, nrpar, ipar, nipar, 0, -1));
    Ranges all;
    all.n = 1; all.b[0] = 0; all.e[0] = neq;
    RS_TRY(fd_jacobian(c, func, w, neq, nnz, ngp, time, d_y.p, group_of.p, rowptr.p, col.p, d_val.p, d_beta.p, all, all));
    RS_TRY(gather_csc_values(c, neq, nnz, d_ian.p, d_jan.p, rowptr.p, col.p, d_val.p, d_a.p));
    if (beta) RS_TRY(d_beta.download(beta, neq, c.stream));
    if (a) RS_TRY(d_a.download(a, (size_t)nnz, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    return RS_OK;
}

extern "C" int rs_residual(const rs_model_desc *func, const double *initpar, int ninitpar, int neq, double time,
                           const double *y, const double *rpar, int nrpar, const int *ipar, int nipar, double *ydot)
{
    if (!y || !ydot) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    RS_TRY(init_model(func, initpar, ninitpar));
    Ctx c;
    DevBuf<double> d_y, d_f;
    RS_TRY(d_y.alloc(neq)); RS_TRY(d_f.alloc(neq));
    RS_TRY(d_y.upload(y, neq, c.stream));
    RS_TRY(d_f.zero(c.stream));
    FdWork w;
    RS_TRY(fd_setup(c, w, neq, 0, 0, rpar, nrpar, ipar, nipar, 0, -1));
    RS_TRY(residual_only(c, func, w, neq, time, d_y.p, d_f.p, neq));
    RS_TRY(d_f.download(ydot, neq, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    return RS_OK;
}

__global__ void shift_down_kernel(int n, const int *__restrict__ in, int *__restrict__ out)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = in[i] - 1;
}

// sorts the (column, value) pairs of every row by column (the solver's row searches need ascending columns)
__global__ void sort_row_pairs_kernel(int n, const int *__restrict__ rowptr, int *__restrict__ col, double *__restrict__ val)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b = rowptr[i], e = rowptr[i + 1];
    for (int p = b + 1; p < e; p++) {
        const int kc = col[p];
        const double kv = val[p];
        int q = p - 1;
        while (q >= b && col[q] > kc) { col[q + 1] = col[q]; val[q + 1] = val[q]; q--; }
        col[q + 1] = kc; val[q + 1] = kv;
    }
}

// row-wise (0-based) pattern + values of J (transposed=0) or J' (transposed=1) from the column-wise input
static int csr_from_columns(Ctx &c, int neq, int nnz, const int *ian, const int *jan, const double *a, int transposed,
                            DevBuf<int> &rowptr, DevBuf<int> &col, DevBuf<double> &val)
{
    DevBuf<int> d_ian, d_jan;
    DevBuf<double> d_a;
    RS_TRY(d_ian.alloc((size_t)neq + 1)); RS_TRY(d_jan.alloc((size_t)nnz)); RS_TRY(d_a.alloc((size_t)nnz));
    RS_TRY(d_ian.upload(ian, (size_t)neq + 1, c.stream));
    RS_TRY(d_jan.upload(jan, (size_t)nnz, c.stream));
    RS_TRY(d_a.upload(a, (size_t)nnz, c.stream));
    RS_TRY(val.alloc((size_t)nnz));
    if (transposed) {  // the columns of J are the rows of J'
        RS_TRY(rowptr.alloc((size_t)neq + 1)); RS_TRY(col.alloc((size_t)nnz));
        shift_down_kernel<<<grid_for(neq + 1, 256), 256, 0, c.stream>>>(neq + 1, d_ian.p, rowptr.p);
        shift_down_kernel<<<grid_for(nnz, 256), 256, 0, c.stream>>>(nnz, d_jan.p, col.p);
        c.launches += 2;
        RS_CUDA(cudaMemcpyAsync(val.p, d_a.p, sizeof(double) * (size_t)nnz, cudaMemcpyDeviceToDevice, c.stream));
        sort_row_pairs_kernel<<<(neq + 255) / 256, 256, 0, c.stream>>>(neq, rowptr.p, col.p, val.p);
        c.launches++;
    } else {
        RS_TRY(build_csr_pattern(c, neq, nnz, d_ian.p, d_jan.p, rowptr, col));
        RS_TRY(scatter_csc_values(c, neq, nnz, d_ian.p, d_jan.p, rowptr.p, col.p, d_a.p, val.p));
    }
    RS_CUDA(cudaStreamSynchronize(c.stream));
    return RS_OK;
}

extern "C" int rs_spmv(int neq, const int *ian, const int *jan, const double *a, const double *x, double *y)
{
    if (!ian || !jan || !a || !x || !y) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    Ctx c;
    const int nnz = ian[neq] - 1;
    DevBuf<int> rowptr, col;
    DevBuf<double> val, d_x, d_y;
    RS_TRY(csr_from_columns(c, neq, nnz, ian, jan, a, 0, rowptr, col, val));
    RS_TRY(d_x.alloc(neq)); RS_TRY(d_y.alloc(neq));
    RS_TRY(d_x.upload(x, neq, c.stream));
    RS_TRY(spmv_launch(c, neq, nnz, rowptr.p, col.p, val.p, d_x.p, d_y.p, pick_lanes(neq, nnz)));
    RS_TRY(d_y.download(y, neq, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    return RS_OK;
}

extern "C" int rs_linsolve(int neq, const int *ian, const int *jan, const double *a, const double *b, int transposed,
                           int nspec_block, const rs_krylov_control *control, double *x, int *iters, double *relres)
{
    if (!ian || !jan || !a || !b || !x) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    Ctx c;
    const int nnz = ian[neq] - 1;
    DevBuf<int> rowptr, col;
    DevBuf<double> val, d_b, d_x;
    RS_TRY(csr_from_columns(c, neq, nnz, ian, jan, a, transposed, rowptr, col, val));
    RS_TRY(d_b.alloc(neq)); RS_TRY(d_x.alloc(neq));
    RS_TRY(d_b.upload(b, neq, c.stream));
    Krylov k;
    const int nspec = nspec_block > 1 ? nspec_block : 1;
    RS_TRY(krylov_setup(c, k, neq, nnz, rowptr.p, col.p, val.p, control ? control->precond : RS_PRECOND_AUTO, nspec,
                        nspec > 1 ? neq / nspec : neq, control, nullptr, nullptr));
    RS_TRY(krylov_precond_update(c, k));
    int it = 0;
    double rr = 0.0;
    RS_TRY(krylov_solve(c, k, d_b.p, d_x.p, 0.0, &it, &rr));
    RS_TRY(d_x.download(x, neq, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    if (iters) *iters = it;
    if (relres) *relres = rr;
    return RS_OK;
}

extern "C" int rs_linsolve_grid(int neq, const int *ian, const int *jan, const double *a, const double *b, const int *grid,
                                const rs_krylov_control *control, double *x, int *iters, double *relres, int *used)
{
    if (!ian || !jan || !a || !b || !x || !grid) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    Ctx c;
    const int nnz = ian[neq] - 1;
    GridDesc g;
    g.type = grid[0]; g.nspec = grid[1];
    const int nd = g.type == RS_TYPE_1D ? 1 : g.type == RS_TYPE_2D ? 2 : g.type == RS_TYPE_3D ? 3 : 0;
    if (!nd) { set_error("grid[0] must be RS_TYPE_1D, 2D or 3D"); return RS_EINVAL; }
    long long cells = 1;
    for (int q = 0; q < nd; q++) { g.d[q] = grid[2 + q]; g.cyc[q] = grid[5 + q]; cells *= g.d[q]; }
    g.ncell = (int)cells;
    if ((long long)g.nspec * cells != neq) { set_error("grid does not match neq"); return RS_EINVAL; }
    DevBuf<int> rowptr, col;
    DevBuf<double> val, d_b, d_x;
    RS_TRY(csr_from_columns(c, neq, nnz, ian, jan, a, 0, rowptr, col, val));
    RS_TRY(d_b.alloc(neq)); RS_TRY(d_x.alloc(neq));
    RS_TRY(d_b.upload(b, neq, c.stream));
    Krylov k;
    RS_TRY(krylov_setup(c, k, neq, nnz, rowptr.p, col.p, val.p, control ? control->precond : RS_PRECOND_AUTO, g.nspec, g.ncell,
                        control, nullptr, nullptr, &g));
    RS_TRY(krylov_precond_update(c, k));
    int it = 0;
    double rr = 0.0;
    RS_TRY(krylov_solve(c, k, d_b.p, d_x.p, 0.0, &it, &rr));
    RS_TRY(d_x.download(x, neq, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    if (iters) *iters = it;
    if (relres) *relres = rr;
    if (used) { used[0] = k.mg_on ? RS_PRECOND_MG : k.btd_on ? RS_PRECOND_BTD : k.precond; used[1] = k.use_cg ? 1 : 0; }
    return RS_OK;
}

extern "C" int rs_errset(int neq, int itol, const double *rtol, const double *atol, const double *y, double *ewt)
{
    if (!rtol || !atol || !y || !ewt) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    Ctx c;
    const int nr = (itol == 3 || itol == 4) ? neq : 1, na = (itol == 2 || itol == 4) ? neq : 1;
    DevBuf<double> d_r, d_a, d_y, d_e;
    RS_TRY(d_r.alloc(nr)); RS_TRY(d_a.alloc(na)); RS_TRY(d_y.alloc(neq)); RS_TRY(d_e.alloc(neq));
    RS_TRY(d_r.upload(rtol, nr, c.stream)); RS_TRY(d_a.upload(atol, na, c.stream)); RS_TRY(d_y.upload(y, neq, c.stream));
    errset_kernel<<<grid_for(neq, 256), 256, 0, c.stream>>>(0, neq, itol, d_r.p, d_a.p, d_y.p, d_e.p);
    RS_TRY(d_e.download(ewt, neq, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    return RS_OK;
}

extern "C" int rs_convergence_test(int neq, const double *beta, const double *ewt, double *out2)
{
    if (!beta || !ewt || !out2 || neq <= 0) { set_error("null argument"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    Ctx c;
    DevBuf<double> d_b, d_e, partials, red;
    DevBuf<unsigned> ticket;
    RS_TRY(d_b.alloc(neq)); RS_TRY(d_e.alloc(neq)); RS_TRY(partials.alloc(2 * 148 * 8)); RS_TRY(red.alloc(2)); RS_TRY(ticket.alloc(1));
    RS_TRY(d_b.upload(beta, neq, c.stream)); RS_TRY(d_e.upload(ewt, neq, c.stream));
    RS_TRY(ticket.zero(c.stream));
    test1_kernel<<<grid_for(neq, 256, 148 * 8), 256, 0, c.stream>>>(0, neq, d_b.p, d_e.p, partials.p, ticket.p, red.p, 0);
    RS_CUDA(cudaGetLastError());
    RS_TRY(red.download(out2, 2, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    return RS_OK;
}

extern "C" const char *rs_last_error(void) { return g_err; }
extern "C" int rs_abi_version(void) { return RS_ABI_VERSION; }
extern "C" int rs_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
extern "C" long long rs_release_cached_memory(void)
{
    {   // the pinned host blocks too (not counted: the return value is device memory)
        PinnedPool &H = pinned_pool();
        std::lock_guard<std::mutex> lk(H.mu);
        for (void *q : H.free_blocks) cudaFreeHost(q);
        H.free_blocks.clear();
    }
    Pool &P = pool();
    std::lock_guard<std::mutex> lk(P.mu);
    return (long long)P.drop_all_locked();
}
