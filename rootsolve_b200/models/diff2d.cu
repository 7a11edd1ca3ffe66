// Bundled model: 2-D diffusion with first-order or Monod consumption and imposed
// boundary values (man/steady.2D.Rd:110-125; BASELINE configs C1 and C3).
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_D2_NPAR)

__device__ void diff2d(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], variant = ip[7];
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const int first = j0 * d2, count = (j1 - j0) * d2;
    const int nthreads = (int)RS_THREAD_COUNT;
    for (int q = (int)RS_THREAD_ID; q < count; q += nthreads) {
        const int idx = first + q;
        const int j = idx / d2, k = idx - j * d2;
        ydot[idx] = rsm_diff2d_point(y, j, k, d1, d2, parms, variant);
    }
}

RS_MODEL(diff2d, RSM_D2_NPAR)
