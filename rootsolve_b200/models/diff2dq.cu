// Bundled model: the vignette's 2-D example (vignettes/rootSolve.Rnw:615-663) -- zero-flux boundaries, quadratic
// consumption, production in a few cells listed in ipar.
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_D2Q_NPAR)

__device__ void diff2dq(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], nsrc = ip[7];
    const int *src = ip + 8;
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    RS_FOR_EACH_CELL(it, k, j0, j1, 1, d2)
        ydot[it.j * d2 + k] = rsm_diff2dq_point(y, it.j, k, d1, d2, parms, src, nsrc);
}

RS_MODEL_ACC(diff2dq, RSM_D2Q_NPAR)
