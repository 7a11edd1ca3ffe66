// Bundled model: the vignette's 2-D example (vignettes/rootSolve.Rnw:615-663) -- zero-flux boundaries, quadratic
// consumption, production in a few cells listed in ipar.
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_D2Q_NPAR)

__device__ void diff2dq(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], nsrc = ip[7];
    const int *src = ip + 8;
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const int first = j0 * d2, count = (j1 - j0) * d2;
    const int nthreads = (int)RS_THREAD_COUNT;
    for (int q = (int)RS_THREAD_ID; q < count; q += nthreads) {
        const int idx = first + q;
        const int j = idx / d2, k = idx - j * d2;
        ydot[idx] = rsm_diff2dq_point(y, j, k, d1, d2, parms, src, nsrc);
    }
}

RS_MODEL(diff2dq, RSM_D2Q_NPAR)
