// Bundled model: two-species reaction-diffusion on a 3-D grid (BASELINE config C4).
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_RD3_NPAR)

__device__ void rd3d(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const int plane = d2 * d3, nc = plane * d1;
    const int first = j0 * plane, count = (j1 - j0) * plane;
    const int nthreads = (int)RS_THREAD_COUNT;
    for (int q = (int)RS_THREAD_ID; q < 2 * count; q += nthreads) {
        const int s = (q >= count) ? 1 : 0;
        const int cell = first + (q - s * count);
        const int j = cell / plane;
        const int rem = cell - j * plane;
        const int k = rem / d3, l = rem - k * d3;
        ydot[s * nc + cell] = rsm_rd3d_point(y, s, j, k, l, d1, d2, d3, parms);
    }
}

RS_MODEL(rd3d, RSM_RD3_NPAR)
