/*
 * example_steady3d.c -- the path called from plain C, the way a compiled host (the R shim next to this file, a
 * Fortran or C++ model driver) would call it: steady.3D of the man page's diffusion model (man/steady.3D.Rd:105-140)
 * on an n x n x n grid, arguments built exactly as steady.3D -> stodes builds them (R/steady.3D.R:44-50,
 * R/stodes.R:132-155).
 *
 *   gcc -std=c99 -Iinclude integration/example_steady3d.c -Lrootsolve_b200 -lrootsolve_b200 -lrs_models \
 *       -Wl,-rpath,$PWD/rootsolve_b200 -o example_steady3d && ./example_steady3d 32
 *
 * Prints nnz, ngp, the precis sequence and whether steady state was reached; exit code 0 on success.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rootsolve_b200.h"

/* the model library exports one descriptor getter per model (RS_MODEL in rootsolve_b200_model.cuh) */
extern const struct rs_model_desc *rs_model_diff3d(void);

int main(int argc, char **argv)
{
    const int n = argc > 1 ? atoi(argv[1]) : 32;
    const int neq = n * n * n;
    const double parms[8] = {1, 1, 1, 1, 1, 1, 0.025, 1};     /* Dx Dy Dz dx dy dz r BND */
    const double tol = 1e-10;
    const int ipar[5] = {1, n, n, n, 0};                       /* what the bundled model reads: nspec, dims, variant */
    int nnz[8] = {0, 1, n, n, n, 0, 0, 0};                     /* (estimate, nspec, rev(dimens), rev(cyclic)) */
    const int positive[1] = {1};
    double *y = (double *)malloc(sizeof(double) * (size_t)neq);
    double *yout = (double *)malloc(sizeof(double) * (size_t)neq);
    double precis[100];
    rs_stsparse_args a;
    rs_stsparse_result r;
    int i, rc;
    if (!y || !yout || n < 1) return 2;
    for (i = 0; i < neq; i++) y[i] = 10.0;
    nnz[0] = neq * (6 + 1) - 3 * 1 * (3 * n);                  /* R/stodes.R:139 */

    memset(&a, 0, sizeof a);
    memset(&r, 0, sizeof r);
    a.y = y; a.neq = neq; a.time = 0.0;
    a.func = rs_model_diff3d();
    a.initpar = parms; a.ninitpar = 8;
    a.ctol = tol; a.atol = &tol; a.natol = 1; a.rtol = &tol; a.nrtol = 1; a.itol = 1;
    a.nnz = nnz; a.nnnz = 8;
    a.lrw = 0; a.ngp = neq + 1; a.maxiter = 100;
    a.Pos = 1; a.positive = positive; a.npositive = 1;
    a.ipar = ipar; a.nipar = 5;
    a.Type = RS_TYPE_3D; a.imethod = 1;
    r.y = yout; r.precis = precis;

    rc = rs_stsparse(&a, &r);
    if (rc != RS_OK) {
        fprintf(stderr, "rs_stsparse failed (%d): %s\n", rc, rs_last_error());
        return 1;
    }
    printf("n=%d nnz=%d ngp=%d steady=%d newton=%d krylov=%d precis=", n, r.steady[1], r.steady[2], r.steady[0], r.niter,
           r.krylov_iters);
    for (i = 0; i < r.niter; i++) printf("%s%.6e", i ? "," : "", precis[i]);
    printf(" y[centre]=%.12g\n", yout[(n / 2) * n * n + (n / 2) * n + n / 2]);
    free(y); free(yout);
    return r.steady[0] == 1 ? 0 : 3;
}
