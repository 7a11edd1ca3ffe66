/*
 * rootsolve_b200.h -- C ABI of the B200-native stodes path.
 *
 * This is the drop-in boundary for rootSolve's sparse-Newton steady-state
 * solver: what R's
 *     .Call("call_stsparse", y, time, Func, initpar, Forc, ctol, atol, rtol, itol,
 *           rho, ModelInit, ModelForc, verbose, nnz, lrw, ngp, maxiter, Pos,
 *           positive, Nglobal, rpar, ipar, Type, ian, jan, imethod, control)
 * (R/stodes.R:330-338; C prototype src/call_stsparse.c:60-64; registered with
 * arity 27 in src/R_init_rootSolve.c:36-41) binds to.  Plain pointers and
 * sizes only; no R, torch or C++ types cross it.  INTEGRATION.md shows the
 * SEXP shim a rootSolve maintainer would put in front of rs_stsparse().
 *
 * Every entry point returns 0 on success or a negative RS_E* code and leaves a
 * message retrievable with rs_last_error() (the reference raises R errors with
 * rexit()/Rf_error(); nothing longjmps across this boundary).  Non-convergence
 * is not an error: steady=0, exactly as src/sparse.f:191-192 +
 * R/stodes.R:371-374.
 *
 * All kernels run on the current CUDA device; there is no CPU fallback -- on a
 * machine without a usable GPU every compute entry point fails with
 * RS_ENODEVICE.
 */
#ifndef ROOTSOLVE_B200_H
#define ROOTSOLVE_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define RS_ABI_VERSION 1

enum {
    RS_OK = 0,
    RS_EINVAL = -1,        /* bad argument (the R front end's stop() checks) */
    RS_ENODEVICE = -2,     /* no CUDA device / CUDA runtime failure */
    RS_ENOMEM = -3,        /* device allocation failed */
    RS_ESTRUCT = -4,       /* "cannot generate sparse jacobian ..." (src/sparsecoeff.f:37-40,193-196) */
    RS_EGROUP = -5,        /* "error during grouping: NGP too small" (src/sparse.f:603-611) */
    RS_EDUPLICATE = -6,    /* "sparse solver: duplicate entry in a" (src/sparse.f:1916,1934) */
    RS_ESINGULAR = -7,     /* zero pivot / Krylov breakdown ("sparse solver: zero pivot", src/sparse.f:2376) */
    RS_EMODEL = -8,        /* model descriptor missing or parameter-length mismatch (src/steady_utils.c:48-53) */
    RS_ECOMM = -9,         /* NCCL failure */
    RS_ETOOBIG = -10       /* size beyond a stated fixed limit */
};

/* Sparsity types, the Type argument of R/stodes.R:66-158 / src/sparse.f:343-353.
 * A negative value means "return the structure only" (sparsetype="sparsereturn"). */
enum { RS_TYPE_USER = 0, RS_TYPE_DETECT = 1, RS_TYPE_1D = 2, RS_TYPE_2D = 3, RS_TYPE_3D = 4,
       /* 2-D / 3-D grid of which only the states listed in a map are unknowns (sparsetype "2Dmap"/"3Dmap",
        * R/stodes.R:111-141, src/testMap.f); the nnz vector then continues after the cyclic flags with TotN =
        * nspec*prod(dims) and the TotN map entries (src/call_stsparse.c:196-206) */
       RS_TYPE_2DMAP = 30, RS_TYPE_3DMAP = 40 };

/* Linear-solver selection: the reference's imethod (1 = yale, 2 = ilut, 3 = ilutp,
 * 10 = yale + return ian/jan; R/stodes.R:314-329).  On the GPU every one of them is
 * served by the preconditioned Krylov solver below (the reference's ilut/ilutp apply an
 * incomplete LU once per Newton step, src/sparsekit.f:141-166: same tests, same steady
 * state, more Newton steps); the ILUT options of `control' have nothing to act on. */
enum { RS_PRECOND_NONE = 0, RS_PRECOND_JACOBI = 1, RS_PRECOND_BLOCK_JACOBI = 2,
       /* aggregation multigrid V-cycle on the 2-D / 3-D grid (slab-local when sharded): what makes the step count of
        * the solve independent of the grid size, the role of the exact factors of nsfc/nnfc (src/sparse.f:1937-2447) */
       RS_PRECOND_MG = 3,
       /* exact solve of the (block-)tridiagonal part by cyclic reduction: 1-D grids (blocks of nspec x nspec) and
        * general structures whose weight sits next to the diagonal */
       RS_PRECOND_BTD = 4,
       RS_PRECOND_AUTO = -1 };

struct rs_model_desc;  /* device residual plug-in, see rootsolve_b200_model.cuh */

/* Krylov controls: the GPU counterpart of stodes' control= list
 * (R/stodes.R:407-420 holds the ILUT knobs of the reference).  Zero / negative
 * fields take the defaults in parentheses. */
typedef struct {
    int precond;            /* RS_PRECOND_* (AUTO: multigrid on 2-D/3-D grids, block-tridiagonal on 1-D grids and banded
                             * general structures, else block-Jacobi on multi-species grids, else Jacobi) */
    int max_krylov;         /* per Newton step (10000) */
    double lin_rtol;        /* ||J x - beta||_2 <= lin_rtol * ||beta||_2 (1e-8: the accuracy of the FD Jacobian) */
    double lin_ewt_factor;  /* ... or rms(J x - beta) <= lin_ewt_factor * min(atol) (5e-4): far below the error
                             * weights, because the error of the state is up to ||J^-1|| times this residual */
    int check_every;        /* host convergence check period in Krylov iterations (4) */
} rs_krylov_control;

/* POD mirror of the 27 .Call arguments (src/call_stsparse.c:60-64). */
typedef struct {
    /* y, time */
    const double *y; int neq; double time;
    /* Func: the compiled residual (R passes a NativeSymbol address, R/stodes.R:212-220) */
    const struct rs_model_desc *func;
    /* initpar / ModelInit: parameters handed to the model's initialiser (src/call_stsparse.c:231-240) */
    const double *initpar; int ninitpar;
    /* Forc / ModelForc */
    const double *forcs; int nforcs;
    /* ctol, atol, rtol, itol (R/stodes.R:56-59) */
    double ctol; const double *atol; int natol; const double *rtol; int nrtol; int itol;
    int verbose;
    /* nnz vector: (nnz_estimate, nspec, dims..., cyclic...) exactly as R/stodes.R:101-155 builds it and
     * src/call_stsparse.c:189-195 reads it */
    const int *nnz; int nnnz;
    int lrw;                /* accepted for signature parity; the Krylov path needs no Yale work array */
    int ngp;                /* MAXG */
    int maxiter;
    int Pos;                /* as.integer(Pos) */
    const int *positive; int npositive;   /* as.integer(positive) */
    int nout;               /* Nglobal */
    const double *rpar; int nrpar;
    const int *ipar; int nipar;
    int Type;
    const int *ian; const int *jan;       /* Type 0 only: user structure, 1-based */
    int imethod;            /* 1, 2, 3 or 10 */
    const rs_krylov_control *control;     /* may be NULL */
} rs_stsparse_args;

/* What call_stsparse returns (src/call_stsparse.c:278-317): yout plus the
 * "precis" and "steady" attributes (and "ian"/"jan" for imethod 10).  The
 * caller owns every buffer; NULL optional buffers are skipped. */
typedef struct {
    double *y;              /* neq + nout */
    double *precis;         /* maxiter doubles; first niter valid */
    int niter;
    int steady[4];          /* isSteady, nnz, ngp, work length (device bytes / 8, or -1 -> NA) */
    int *ian;               /* optional, neq+1, 1-based */
    int *jan;               /* optional, capacity nnz[0], 1-based, the order sparsereturn gives */
    int *igp;               /* optional, ngp(MAXG)+1 */
    int *jgp;               /* optional, neq */
    /* extras with no counterpart in the reference */
    int krylov_iters;       /* total Krylov iterations */
    int kernel_launches;    /* kernels launched by this call */
    double t_setup_s;       /* structure + grouping + CSR build (device events) */
    double t_newton_s;      /* Newton loop */
    int spmv_kernel;        /* SpMV the Krylov solver ran: 0 sub-warp CSR, 1 TMA-staged CSR, 2 TMA-staged with
                             * implicit columns (non-cyclic grid, <= 2 species) */
    int precond;            /* preconditioner the last linear solve ran under (RS_PRECOND_*); a handle that had to
                             * drop the multigrid cycle reports the (block-)Jacobi it fell back to */
    int krylov_method;      /* 0 BiCGStab, 1 CG (Jacobian symmetric to finite-difference noise) */
} rs_stsparse_result;

/* The path: replaces call_stsparse -> dsparse (src/call_stsparse.c:60-320, src/sparse.f:30-195). */
int rs_stsparse(const rs_stsparse_args *args, rs_stsparse_result *res);

/* ---- device-resident handle (same path, inputs/outputs staying in HBM) ---- */
typedef struct rs_solver rs_solver;
/* Builds structure, grouping, CSR and all work vectors once (src/sparse.f:89-105 equivalent). */
int rs_solver_create(const rs_stsparse_args *args, rs_solver **out);
/* Copies a new initial state into the handle from host (is_device=0) or device (1) memory. */
int rs_solver_set_state(rs_solver *s, const double *y, int is_device);
/* Runs the Newton loop (src/sparse.f:114-188) from the state in the handle. */
int rs_solver_run(rs_solver *s, rs_stsparse_result *res_meta);
/* Device pointer of the state vector (neq doubles). */
double *rs_solver_state_device(rs_solver *s);
int rs_solver_get_state(rs_solver *s, double *y_host);
/* rs_solver_set_profiling(s, 1) brackets every launch with CUDA events on the solver's stream (and resets
 * the counters); rs_solver_kernel_stats then reports the per-kernel device time since: fills
 * names[i] (static strings), ms[i], launches[i], bytes[i] (algorithmic bytes per
 * launch) for up to cap kernels; returns the count. */
int rs_solver_set_profiling(rs_solver *s, int on);
int rs_solver_kernel_stats(rs_solver *s, int cap, const char **names, double *ms, long long *launches,
                           double *bytes);
void rs_solver_destroy(rs_solver *s);

/* ---- building blocks exposed for parity tests (host buffers in and out) ---- */

/* Replaces sparse1d/sparse2d/sparse3d (src/sparsecoeff.f:5-83, 89-176, 207-312).
 * type = RS_TYPE_1D/2D/3D; dims/cyclic as stodes passes them (rev(dimens), rev(Bnd)).
 * nnz: in = capacity of jan, out = number of nonzeros. */
int rs_sparse_structure(int type, int neq, int nspec, const int *dims, const int *cyclic, int *nnz,
                        int *ian, int *jan);

/* Replaces sparse2dmap/sparse3dmap (src/testMap.f:25-115, 119-227).  type = RS_TYPE_2DMAP/3DMAP; neq = number
 * of states present; pres[totn], totn = nspec*prod(dims): new number (1..neq, in grid order) of each full-grid
 * state, <= 0 if absent.  nnz: in = capacity, out = count; as in the reference the capacity must EXCEED the
 * count (updatejan, src/testMap.f:17-20). */
int rs_sparse_structure_map(int type, int neq, int nspec, const int *dims, const int *cyclic, const int *pres,
                            int totn, int *nnz, int *ian, int *jan);

/* Replaces JGROUP (src/sparse.f:531-616).  grid_hint: 0 = generic algorithm; otherwise
 * {type, nspec, d1,d2,d3, cyc1,cyc2,cyc3} lets the solver use the wavefront variant. */
int rs_jgroup(int neq, const int *ian, const int *jan, int maxg, const int *grid_hint, int *ngp, int *igp,
              int *jgp);

/* Replaces xSparseJacob (src/sparse.f:461-525): beta = -f(y) and a[] in jan order. */
int rs_fd_jacobian(const struct rs_model_desc *func, const double *initpar, int ninitpar, int neq,
                   double time, const double *y, const int *ian, const int *jan, const int *igp,
                   const int *jgp, int ngp, const double *rpar, int nrpar, const int *ipar, int nipar,
                   double *beta, double *a);

/* One residual evaluation f(t, y) through the device model (what XMODEL does, src/sparse.f:488). */
int rs_residual(const struct rs_model_desc *func, const double *initpar, int ninitpar, int neq, double time,
                const double *y, const double *rpar, int nrpar, const int *ipar, int nipar, double *ydot);

/* Stands in for cdrv (src/sparse.f:1427-1697): solves J x = b (transposed=0) or J' x = b (1), J given by
 * columns as (ian, jan, a).  Returns Krylov iterations in *iters, final relative residual in *relres. */
int rs_linsolve(int neq, const int *ian, const int *jan, const double *a, const double *b, int transposed,
                int nspec_block, const rs_krylov_control *control, double *x, int *iters, double *relres);

/* The same solve for a matrix that lives on a 1-D / 2-D / 3-D grid (the structures of sparse1d/2d/3d,
 * src/sparsecoeff.f:5-312), so that the grid preconditioners -- multigrid, block-tridiagonal -- can be exercised
 * on their own.  grid = { RS_TYPE_1D/2D/3D, nspec, d1, d2, d3, cyc1, cyc2, cyc3 } with unused extents 1 (d1 slowest,
 * as in the nnz vector of stodes).  *used = { preconditioner (RS_PRECOND_*), method (0 BiCGStab, 1 CG) } the solve ran. */
int rs_linsolve_grid(int neq, const int *ian, const int *jan, const double *a, const double *b, const int *grid,
                     const rs_krylov_control *control, double *x, int *iters, double *relres, int *used);

/* y = J x with J given by columns as (ian, jan, a) -- the SpMV kernel on its own. */
int rs_spmv(int neq, const int *ian, const int *jan, const double *a, const double *x, double *y);

/* Convergence test 1 of the Newton loop (src/sparse.f:122-133): out2[0] = sum |beta_k| (precis is this / neq),
 * out2[1] = max |beta_k / ewt_k| (steady when <= 1). */
int rs_convergence_test(int neq, const double *beta, const double *ewt, double *out2);

/* No counterpart in the reference.  The same test as rs_convergence_test for callers that cannot allocate
 * device memory while they run (applications that gave the GPU heap to their own pool, in-situ / real-time
 * coupling loops where cudaMalloc would stall the device, plugins barred from allocating after start-up):
 * d_beta, d_ewt (neq doubles each) and d_out2 (2 doubles) are DEVICE pointers the caller owns; the kernel is
 * queued on `stream` (a cudaStream_t, NULL = default stream) and the call returns without synchronising.
 * Scratch comes from a 64 KiB static device array inside the library, hence the limit: neq must not exceed
 * RS_CONVERGENCE_TEST_NOALLOC_MAX_NEQ (RS_ETOOBIG otherwise).  No cudaMalloc/cudaFree is ever called.
 * The array is shared by all calls on a device: calls that may overlap on the GPU race and must be ordered
 * by the caller.  The first call on each device resolves the array's address; make it before any stream
 * capture begins. */
#define RS_CONVERGENCE_TEST_NOALLOC_MAX_NEQ 8355840
int rs_convergence_test_noalloc(int neq, const double *d_beta, const double *d_ewt, double *d_out2, void *stream);

/* errset (src/sparse.f:264-293) */
int rs_errset(int neq, int itol, const double *rtol, const double *atol, const double *y, double *ewt);

const char *rs_last_error(void);
int rs_abi_version(void);
/* number of visible CUDA devices (0 when there is none); never fails */
int rs_device_count(void);
/* Device work arrays are cached between calls (the reference gets the same effect from R_alloc'ing out of R's own
 * heap, src/call_stsparse.c:108-225: no system allocation per call).  Returns the cached bytes to the driver and
 * reports how many there were; RS_POOL_MAX_BYTES in the environment caps the cache (0: none). */
long long rs_release_cached_memory(void);

/* ---- multi-GPU (one process per GPU; slabs over the slowest grid index) ---- */
/* 128-byte NCCL unique id: rank 0 calls rs_comm_unique_id and ships the bytes to the other ranks
 * (the Python host layer does it with torch.distributed); then every rank calls rs_comm_init. */
int rs_comm_unique_id(char id[128]);
int rs_comm_init(int nranks, int rank, const char id[128]);
int rs_comm_nranks(void);
int rs_comm_rank(void);
void rs_comm_finalize(void);
/* slab owned by a rank: planes [*j0, *j1) of the slowest grid dimension d1 */
void rs_slab_bounds(int d1, int nranks, int rank, int *j0, int *j1);

#ifdef __cplusplus
}
#endif
#endif /* ROOTSOLVE_B200_H */
