/*
 * stodes_oracle.h -- CPU oracle for the stodes sparse-Newton steady-state path.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a plain-C restatement of rootSolve's
 * algorithm for the path (src/sparse.f, src/sparsecoeff.f, src/steady.f:208-226,
 * src/call_stsparse.c), used as the checker in tests/, in
 * __graft_entry__.smoke() and as the timed CPU baseline / reference arm in
 * bench.py.  Nothing under rootsolve_b200/ may include, link or call it.
 *
 * Parity status: PINNED against the goldens the reference ships for this path:
 * inst/doc/rootSolve.pdf p.16 (nnz=6400, ngp=12, work length 82366, 3 iterations,
 * all three printed precis values), pdf p.8-9 (the compiled anoxic model: y, SumS,
 * eight precis values), man/steady.2D.Rd:202-206 (the cyclic example needs exactly
 * the lrw = 283800 it is called with) and the closed-form nnz/ngp counts -- see
 * tests/test_oracle_golden.py.  The reference itself (R + Fortran 77) cannot be
 * compiled in this image (no R, no Fortran compiler), so there is no
 * oracle/_ref build; DESIGN.md says so.
 *
 * All index arrays are 1-based exactly as in the Fortran (ian[0]=1 ...).
 */
#ifndef STODES_ORACLE_H
#define STODES_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* residual with the compiled-model signature, src/steady.h:34 */
typedef void (*orc_deriv_t)(int *neq, double *t, double *y, double *ydot, double *yout, int *ip);

/* structure generators, src/sparsecoeff.f:5-83, 89-176, 207-312.
 * nnz: in = capacity of jan, out = number of nonzeros.  Return 0 or <0 error. */
int orc_sparse1d(int ntot, int nspec, int dim, int cyclic, int *nnz, int *ian, int *jan);
int orc_sparse2d(int ntot, int nspec, const int *dimens, const int *cyclic, int *nnz, int *ian, int *jan);
int orc_sparse3d(int ntot, int nspec, const int *dimens, const int *cyclic, int *nnz, int *ian, int *jan);

/* mapped (masked) grids, src/testMap.f:25-115, 119-227.  pres[M-1] = new number of full-grid state M (1..ntot,
 * increasing with M) or <= 0 if the state is absent; totn = nspec * prod(dimens) = length of pres. */
int orc_sparse2dmap(int ntot, int nspec, const int *dimens, const int *cyclic, int *nnz, int *ian, int *jan,
                    const int *pres, int totn);
int orc_sparse3dmap(int ntot, int nspec, const int *dimens, const int *cyclic, int *nnz, int *ian, int *jan,
                    const int *pres, int totn);

/* column grouping, src/sparse.f:531-616.  igp has room for maxg+1 entries.
 * Returns 0, 1 (ier: not all columns chosen) or 2 (more than maxg groups). */
int orc_jgroup(int n, const int *ia, const int *ja, int maxg, int *ngrp, int *igp, int *jgp);

/* FD increment, src/steady.f:208-226 */
double orc_perturb(double *value);

/* error weights, src/sparse.f:264-293 */
void orc_errset(int n, int itol, const double *rtol, const double *atol, const double *y, double *ewt);

/* grouped FD Jacobian, src/sparse.f:461-525.  beta = -f(y), a[] in jan order. */
void orc_sparse_jacob(orc_deriv_t f, int n, const int *ian, const int *jan, const int *igp,
                      const int *jgp, int ngp, double *svar, double *dsvar, double *beta,
                      double t, double *out, int *ipar, double *a);

/* Yale SMP pieces exposed for unit tests */
int orc_odrv(int n, const int *ia, const int *ja, int *p, int *ip, int nsp, int *isp, int path);
int orc_cdrv(int n, int *r, int *c, int *ic, int *ia, int *ja, double *a, double *b, double *z,
             int nsp, int *isp, double *rsp, int *esp, int path);

typedef struct {
    /* inputs (mirror of the 27 .Call arguments that matter off-R, src/call_stsparse.c:60-64) */
    int neq;
    double time;
    orc_deriv_t func;
    double ctol;
    const double *atol; int natol;
    const double *rtol; int nrtol;
    int itol;
    const int *nnzvec; int nnnz;   /* (nnz_est, nspec, dims..., cyclic...) as R/stodes.R:101-155 builds it */
    int lrw;                       /* nsp */
    int ngp;                       /* MAXG */
    int maxiter;
    int posit;                     /* as.integer(Pos) */
    const int *pos; int ipos;      /* as.integer(positive) */
    int nout;
    const double *rpar; int nrpar;
    const int *ipar; int nipar;
    int type;                      /* 0,+-1,+-2,+-3,+-4,+-30,+-40 */
    const int *ian_in; const int *jan_in;   /* type 0 only */
    int imethod;                   /* 1 yale, 10 sparsereturn */
    int quirk_first_step_transposed; /* 1 = as the reference (A.3-1), 0 = always solve J x = beta */
    int quirk_unterminated_igp;      /* 1 = as the reference: igp(ngp+1) left at 0 when ngp == neq, last column of J skipped */
} orc_args;

typedef struct {
    double *y;        /* neq + nout */
    double *precis;   /* maxiter */
    int niter;
    int steady;       /* isSteady */
    int dims[3];      /* nnz, ngp, nsp-esp */
    int *ian;         /* neq+1 (filled always) */
    int *jan;         /* capacity nnz_est (filled always; as returned by sparsereturn, i.e. before nroc) */
    int *igp;         /* maxg+1 */
    int *jgp;         /* neq */
    double *a0;       /* capacity nnz_est: FD Jacobian of the first iteration in jan order, before nroc (may be NULL) */
    double *beta0;    /* neq: -f(y0) (may be NULL) */
    /* per-phase wall-clock seconds */
    double t_struct, t_order, t_jac, t_factor, t_solve, t_total;
    int status;       /* 0 ok, else error */
    char msg[256];
} orc_result;

/* whole solve, src/call_stsparse.c:60-320 + src/sparse.f:30-195 */
int orc_stsparse(const orc_args *args, orc_result *res);

#ifdef __cplusplus
}
#endif
#endif
