// Bundled model: diff3d plus the source term of a manufactured solution (closed-form discrete steady state at any
// grid size; SURVEY.md 8(c), parity protocol step 3).  The three sine tables arrive in rpar.
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_D3MS_NPAR)

__device__ void diff3dms(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    const double *tab = yout + ip[0];
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const int plane = d2 * d3;
    const int first = j0 * plane, count = (j1 - j0) * plane;
    const int nthreads = (int)RS_THREAD_COUNT;
    for (int q = (int)RS_THREAD_ID; q < count; q += nthreads) {
        const int idx = first + q;
        const int j = idx / plane;
        const int rem = idx - j * plane;
        const int k = rem / d3, l = rem - k * d3;
        ydot[idx] = rsm_diff3dms_point(y, j, k, l, d1, d2, d3, parms, tab);
    }
}

RS_MODEL(diff3dms, RSM_D3MS_NPAR)
