/*
 * call_stsparse_b200.c -- the R-side binding a rootSolve maintainer would add to src/ so that stodes() runs on the
 * GPU library: a .Call entry point with the arity and argument order of call_stsparse (src/call_stsparse.c:60-64,
 * registered with 27 arguments in src/R_init_rootSolve.c:36-41).  It only unpacks the SEXPs into the POD mirror
 * rs_stsparse_args, calls rs_stsparse() and repacks the result with the attributes R/stodes.R:340-397 reads
 * ("precis", "steady", and "ian"/"jan" for sparsereturn).
 *
 * Not built in this repository (no R here); tests/test_host_api.py only checks that it is valid C against the
 * minimal declarations in integration/mock_r/.  With R: R CMD SHLIB call_stsparse_b200.c -lrootsolve_b200.
 */
#include <string.h>

#include <R.h>
#include <Rdefines.h>

#include "rootsolve_b200.h"

SEXP call_stsparse_b200(SEXP y, SEXP time, SEXP func, SEXP parms, SEXP forcs, SEXP chtol, SEXP atol, SEXP rtol,
                        SEXP itol, SEXP rho, SEXP initfunc, SEXP initforc, SEXP verbose, SEXP NNZ, SEXP NSP,
                        SEXP MAXG, SEXP nIter, SEXP Posit, SEXP Pos, SEXP nOut, SEXP Rpar, SEXP Ipar, SEXP Type,
                        SEXP Ian, SEXP Jan, SEXP Met, SEXP Option)
{
    rs_stsparse_args a;
    rs_stsparse_result r;
    SEXP yout, P, S, IA, JA;
    int n = LENGTH(y), nout = INTEGER(nOut)[0], maxit = INTEGER(nIter)[0], nprot = 0;
    (void)rho; (void)Option;            /* R closures cannot run on the device; the ILUT options have nothing to act on */
    memset(&a, 0, sizeof a);
    memset(&r, 0, sizeof r);
    a.y = REAL(y); a.neq = n; a.time = REAL(time)[0];
    /* func: external pointer to the model's descriptor getter rs_model_<name>() (see rootsolve_b200_model.cuh) */
    a.func = ((const struct rs_model_desc *(*)(void)) R_ExternalPtrAddrFn(func))();
    if (!isNull(initfunc) && LENGTH(parms) > 0) { a.initpar = REAL(parms); a.ninitpar = LENGTH(parms); }
    if (!isNull(initforc) && LENGTH(forcs) > 0) { a.forcs = REAL(forcs); a.nforcs = LENGTH(forcs); }
    a.ctol = REAL(chtol)[0];
    a.atol = REAL(atol); a.natol = LENGTH(atol);
    a.rtol = REAL(rtol); a.nrtol = LENGTH(rtol);
    a.itol = INTEGER(itol)[0];
    a.verbose = INTEGER(verbose)[0];
    a.nnz = INTEGER(NNZ); a.nnnz = LENGTH(NNZ);
    a.lrw = INTEGER(NSP)[0];
    a.ngp = INTEGER(MAXG)[0];
    a.maxiter = maxit;
    a.Pos = INTEGER(Posit)[0]; a.positive = INTEGER(Pos); a.npositive = LENGTH(Pos);
    a.nout = nout;
    a.rpar = REAL(Rpar); a.nrpar = LENGTH(Rpar);
    a.ipar = INTEGER(Ipar); a.nipar = LENGTH(Ipar);
    a.Type = INTEGER(Type)[0];
    a.ian = INTEGER(Ian); a.jan = INTEGER(Jan);
    a.imethod = INTEGER(Met)[0];

    PROTECT(yout = allocVector(REALSXP, n + nout)); nprot++;
    r.y = REAL(yout);
    r.precis = (double *) R_alloc(maxit, sizeof(double));
    if (a.imethod == 10) {
        r.ian = (int *) R_alloc(n + 1, sizeof(int));
        r.jan = (int *) R_alloc(INTEGER(NNZ)[0], sizeof(int));
    }
    if (rs_stsparse(&a, &r) != RS_OK) error("%s", rs_last_error());     /* what rexit() does in the Fortran */

    PROTECT(P = allocVector(REALSXP, r.niter)); nprot++;
    memcpy(REAL(P), r.precis, r.niter * sizeof(double));
    setAttrib(yout, install("precis"), P);
    PROTECT(S = allocVector(INTSXP, 4)); nprot++;
    memcpy(INTEGER(S), r.steady, 4 * sizeof(int));
    setAttrib(yout, install("steady"), S);
    if (a.imethod == 10) {                                              /* src/call_stsparse.c:299-317 */
        PROTECT(IA = allocVector(INTSXP, n + 1)); nprot++;
        memcpy(INTEGER(IA), r.ian, (n + 1) * sizeof(int));
        setAttrib(yout, install("ian"), IA);
        PROTECT(JA = allocVector(INTSXP, r.steady[1])); nprot++;
        memcpy(INTEGER(JA), r.jan, r.steady[1] * sizeof(int));
        setAttrib(yout, install("jan"), JA);
    }
    UNPROTECT(nprot);
    return yout;
}
