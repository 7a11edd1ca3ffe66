// Bundled model: 3-D diffusion with first-order consumption and imposed boundary
// values (man/steady.3D.Rd:105-140, vignettes/rootSolve.Rnw:686-739; BASELINE
// config C5 and the 10^3 golden case).
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_D3_NPAR)

__device__ void diff3d(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6], variant = ip[7];
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    // 32-bit index arithmetic: neq < 2^31, and 64-bit divisions would dominate this kernel
    const int plane = d2 * d3;
    const int first = j0 * plane, count = (j1 - j0) * plane;
    const int nthreads = (int)RS_THREAD_COUNT;
    for (int q = (int)RS_THREAD_ID; q < count; q += nthreads) {
        const int idx = first + q;
        const int j = idx / plane;
        const int rem = idx - j * plane;
        const int k = rem / d3, l = rem - k * d3;
        ydot[idx] = rsm_diff3d_point(y, j, k, l, d1, d2, d3, parms, variant);
    }
}

RS_MODEL(diff3d, RSM_D3_NPAR)
