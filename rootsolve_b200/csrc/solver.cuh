// solver.cuh -- internal interfaces between the translation units of the library.
#pragma once

#include "rs_internal.cuh"

namespace rs {

// Regular-grid description (what dims[] carries in src/sparse.f:398-434).
struct GridDesc {
    int type = 0;          // RS_TYPE_1D/2D/3D, 0 = no grid knowledge
    int nspec = 1;
    int d[3] = {1, 1, 1};  // rev(dimens): d[0] slowest
    int cyc[3] = {0, 0, 0};
    int ncell = 0;
};

// The unknowns a rank works on: up to MAXR disjoint index ranges (one per species of its slab;
// a single range [0, N) on one GPU).  Vectors keep GLOBAL indices on every rank.
constexpr int MAXR = 16;
struct Ranges {
    int n = 0;
    int b[MAXR], e[MAXR];
    long long total() const { long long t = 0; for (int q = 0; q < n; q++) t += e[q] - b[q]; return t; }
};

// One halo exchange: plane j0 / j1-1 of every species go to the slab neighbours, planes j0-1 / j1 come back.
struct HaloPlan {
    int nspec = 0, ncell = 0, plane = 0, j0 = 0, j1 = 0;
    bool has_lo = false, has_hi = false;
};

// structure.cu
int exclusive_scan(Ctx &c, const int *in, int *out, int n, int base, bool write_total);
int build_grid_structure(Ctx &c, const GridDesc &g, int n, DevBuf<int> &d_ian, DevBuf<int> &d_jan, int *nnz);
int build_mapped_structure(Ctx &c, const GridDesc &g, int n, int totn, const int *h_pres, DevBuf<int> &d_ian,
                           DevBuf<int> &d_jan, int *nnz);
int build_csr_pattern(Ctx &c, int n, int nnz, const int *d_ian, const int *d_jan, DevBuf<int> &rowptr,
                      DevBuf<int> &col);
// non-cyclic 1-D / 2-D / 3-D grids: the same pattern written down in closed form (structure.cu)
bool grid_pattern_applicable(const GridDesc &g);
int build_csr_pattern_grid(Ctx &c, const GridDesc &g, int n, int nnz, const int *d_ian, const int *d_jan, DevBuf<int> &rowptr,
                           DevBuf<int> &col);

// jgroup.cu -- group_of[j] = 0-based group of column j; ngp = number of groups.
int jgroup_device(Ctx &c, const GridDesc &g, int n, int nnz, const int *d_ian, const int *d_jan,
                  const int *d_rowptr, const int *d_col, int maxg, DevBuf<int> &group_of, int *ngp);
// igp/jgp in the reference's layout (1-based) from group_of
int jgroup_lists(Ctx &c, int n, int ngp, const int *d_group_of, int *igp_host, int *jgp_host);

// fdjac.cu
struct FdWork {
    DevBuf<double> Y, F;     // (ngp+1) x ld
    DevBuf<double> yout;     // (ngp+1) x ldout
    DevBuf<int> ip;          // reference ipar block + slab bounds
    long long ld = 0, ldout = 0;
    int ngp_alloc = 0;
    // non-cyclic grid with 1-2 species whose row-wise pattern krylov_setup verified (0 = general structure): the
    // assembly then computes the column of every entry from the row's grid position instead of reading col[]
    int gridcols = 0, gd1 = 1, gd2 = 1, gd3 = 1, gncell = 0;
};
int fd_setup(Ctx &c, FdWork &w, int n, int ngp, int nout, const double *rpar, int nrpar, const int *ipar,
             int nipar, int slab_j0, int slab_j1);
// beta = -f(y); val[] (row-wise) = FD Jacobian.  `own`: rows this rank assembles; `touch`: the
// unknowns whose perturbed copies it needs (own + halo planes).
int fd_jacobian(Ctx &c, const rs_model_desc *m, FdWork &w, int n, int nnz, int ngp, double t, const double *d_y,
                const int *d_group_of, const int *d_rowptr, const int *d_col, double *d_val, double *d_beta,
                const Ranges &own, const Ranges &touch);
// the same in the two halves the Newton loop uses: beta = -f(y) first (enough for convergence test 1), then the
// ngp perturbed evaluations + assembly only when a step is actually taken
int fd_base(Ctx &c, const rs_model_desc *m, FdWork &w, int n, double t, const double *d_y, double *d_beta, const Ranges &own);
int fd_groups(Ctx &c, const rs_model_desc *m, FdWork &w, int n, int nnz, int ngp, double t, const double *d_y,
              const int *d_group_of, const int *d_rowptr, const int *d_col, double *d_val, double *d_beta,
              const Ranges &own, const Ranges &touch);
int residual_only(Ctx &c, const rs_model_desc *m, FdWork &w, int n, double t, const double *d_y, double *d_f,
                  long long work);
// numeric structure detection (Type +-1): ian/jan from N perturbed evaluations, bit-exact with src/sparse.f:356-396
int detect_structure(Ctx &c, const rs_model_desc *m, int n, double t, const double *d_y, int nout, const double *rpar,
                     int nrpar, const int *ipar, int nipar, int nnz_cap, DevBuf<int> &ian, DevBuf<int> &jan, int *nnz);
// a_csc[k] = val[position of (jan[k], column of k)] -- for returning the Jacobian in jan order
int gather_csc_values(Ctx &c, int n, int nnz, const int *d_ian, const int *d_jan, const int *d_rowptr,
                      const int *d_col, const double *d_val, double *d_a);
int scatter_csc_values(Ctx &c, int n, int nnz, const int *d_ian, const int *d_jan, const int *d_rowptr,
                       const int *d_col, const double *d_a, double *d_val);

// mg.cu -- aggregation multigrid preconditioner for 2-D / 3-D grids (slab-local when sharded)
struct Mg;
bool mg_applicable(const GridDesc &g, int j0, int j1);
int mg_setup(Ctx &c, Mg **out, const GridDesc &g, int j0, int j1);   // planes [j0, j1) of the slowest dimension; j1 <= j0: all
// grid_order: the rows were verified to be the grid's own stencil in ascending column order (grid_stencil.cuh)
int mg_update(Ctx &c, Mg &m, const int *rowptr, const int *col, const double *val, bool grid_order);
// z = M^-1 r (one V-cycle) on the rank's slab of global-index vectors; dot_out: <r, z>, all-reduced through pr
int mg_apply(Ctx &c, Mg &m, const double *r, double *z, double *partials, unsigned *ticket, double *dot_out, const PeerRed &pr);
int mg_levels(const Mg &m);
// { max |a_ij - a_ji|, max |a_ij| } (device) read off the level-0 copy of the matrix mg_update was last given
int mg_asymmetry(Ctx &c, Mg &m, double *partials, unsigned *ticket, double *out2);
void mg_destroy(Mg *m);

// btd.cu -- block-tridiagonal part solved exactly by cyclic reduction (1-D grids, banded general structures)
struct Btd;
int btd_setup(Ctx &c, Btd **out, int ns, int nb, int stride, int off);
int btd_update(Ctx &c, Btd &t, const int *rowptr, const int *col, const double *val);
int btd_apply(Ctx &c, Btd &t, const double *r, double *z);
int btd_band_share(Ctx &c, int n, const int *rowptr, const int *col, const double *val, double *partials, unsigned *ticket,
                   double *d_out2, double *share);
void btd_destroy(Btd *t);

// krylov.cu
struct Krylov {
    int n = 0, nnz = 0;
    const int *rowptr = nullptr, *col = nullptr;
    const double *val = nullptr;
    int precond = RS_PRECOND_JACOBI, nspec = 1, ncell = 0;
    Ranges own;                 // rows of this rank
    int cell0 = 0, cell1 = 0;   // cells of this rank (block preconditioner)
    HaloPlan halo;
    DevBuf<double> r, r0, p, v, s, t, ph, sh, dinv, binv, z;
    Mg *mg = nullptr;          // multigrid preconditioner; z = M^-1 r of preconditioned CG
    bool mg_on = false;        // the current solve is preconditioned by the cycle
    bool mg_failed = false;    // a solve with it did not converge: this handle goes back to (block-)Jacobi
    bool diag_valid = false;   // dinv / binv hold the current matrix (formed lazily while the cycle or the band solve is in charge)
    Btd *btd = nullptr;        // block-tridiagonal preconditioner (same life cycle as mg)
    bool btd_on = false, btd_failed = false;
    int pool_generation = 0;   // of the peer-mapped buffers this handle points into (sharded solves)
    int btd_decide = 0;        // general structure: 1 = decide at the first update from the band's share of the weight
    DevBuf<double> partials, scal;
    DevBuf<unsigned> ticket;
    double *h_scal = nullptr;  // pinned mirror
    int lanes = 8;             // lanes per row of the sub-warp SpMV
    int tma_ok = 0, tma_sms = 0, tma_per_sm = 0;  // the TMA-staged SpMV is usable (every tile fits a stage)
    // Non-cyclic 1-D/2-D/3-D grid with 1 or 2 species whose row-wise pattern was verified against the stencil:
    // the SpMV then COMPUTES the column of every entry from the row's grid position and never reads col[]
    // (gridcols = number of species, 0 = general CSR).  gd1..gd3: the grid as a 3-D box (leading extents 1).
    int gridcols = 0, gd1 = 1, gd2 = 1, gd3 = 1, tma_per_sm_grid = 0;
    double *ph_v = nullptr, *sh_v = nullptr;   // the preconditioned directions (SpMV inputs): own buffers or pooled, peer-mapped ones
    bool use_cg = false;                   // J is symmetric to FD noise: preconditioned CG (one SpMV per step) instead of BiCGStab
    double asym = 0.0;                     // max |a_ij - a_ji| / max |a_ij| of the last Jacobian
    // CG on the symmetrically Jacobi-scaled system sigma S A S (values scaled in place): s, the common sign of the
    // diagonal, and the factor that turns the recurrence's ||r^|| into a residual norm of the original system
    DevBuf<double> sc;
    bool scaled = false;
    double sigma = 1.0, norm_conv = 1.0;
    bool g_scaled[2] = {false, false}, g_mg[2] = {false, false};
    bool push_ok = false;                  // halos of ph / sh are pushed into the neighbours' vectors over NVLink
    void *nb_lo[2] = {nullptr, nullptr}, *nb_hi[2] = {nullptr, nullptr};   // neighbours' ph (0) and sh (1)
    // captured Krylov steps ([0] BiCGStab, [1] CG): graph, the solution vector it updates, steps and launches per replay
    cudaGraphExec_t g_exec[2] = {nullptr, nullptr};
    const double *g_x[2] = {nullptr, nullptr};
    int g_n[2] = {0, 0}, g_launches[2] = {0, 0};
    // The host looks at ||r|| every check_every steps: each look drains the stream (a ~20 us bubble, as
    // long as a whole step on 8 GPUs or on a small grid).  Steps taken after convergence are harmless --
    // a breakdown there leaves x untouched (update_xr_kernel) -- and cost at most check_every-1 steps.
    int max_iter = 10000, check_every = 4;
    // lin_rtol: the finite-difference Jacobian carries a relative error of a few 1e-8 (Perturb's increment), so a
    // Newton step cannot gain from a linear solve more accurate than that; the residual left over is picked up by
    // the next step, whose own solve stops at the absolute floor (lin_ewt_factor, see rs_solver_run).
    double lin_rtol = 1e-8, lin_ewt_factor = 5e-4;
    ~Krylov();
};
int krylov_setup(Ctx &c, Krylov &k, int n, int nnz, const int *rowptr, const int *col, const double *val,
                 int precond, int nspec, int ncell, const rs_krylov_control *ctl, const Ranges *own,
                 const HaloPlan *halo, const GridDesc *grid = nullptr);
int krylov_precond_update(Ctx &c, Krylov &k);
// solves A x = b (x overwritten, starts from 0); abs_floor: absolute 2-norm stop.  *converged: the true residual
// reached the solve's own target max(lin_rtol ||b||, abs_floor) -- which can leave relres above any fixed fraction
// when ||b|| is already close to the floor
int krylov_solve(Ctx &c, Krylov &k, const double *b, double *x, double abs_floor, int *iters, double *relres,
                 bool *converged = nullptr);
int spmv_launch(Ctx &c, int n, int nnz, const int *rowptr, const int *col, const double *val, const double *x,
                double *y, int lanes);
int pick_lanes(int n, int nnz);

// comm.cu
bool comm_active();
int comm_nranks();
int comm_rank();
// NVLink peer-memory all-reduce fused into the reducing kernels (rs_internal.cuh); falls back to NCCL when off
bool comm_peer_enabled();
PeerRed comm_next_peer_red();   // descriptor with the next sequence number (nranks = 1 when disabled)
// halo push: map the slab neighbours' copies of vectors (collective); handshake flags with the next sequence number
bool comm_share_vectors(void *const *mine, int count, void **lo, void **hi);
void comm_unshare_vectors(int count, void **lo, void **hi);
// two n-vectors from a pool that stays mapped into the slab neighbours for the life of the communicator
bool comm_pooled_vectors(size_t n, double **mine, void **lo, void **hi);
// one more peer-mapped buffer of n doubles (multigrid level vectors whose halo planes travel); same life cycle
bool comm_pooled_arena(size_t n, double **mine, void **lo, void **hi);
int comm_pool_generation();   // changes whenever those buffers are re-created (a larger solve asked for more)
HaloSignal comm_next_halo_signal();
int comm_allreduce_sum(Ctx &c, double *d_buf, int count);
int comm_allreduce_max(Ctx &c, double *d_buf, int count);
int comm_halo_exchange(Ctx &c, double *v, const HaloPlan &h);
int comm_gather_slabs(Ctx &c, double *v, int nspec, int ncell, int plane, int d1);

}  // namespace rs
