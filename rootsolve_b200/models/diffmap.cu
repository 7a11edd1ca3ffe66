// Bundled model: diffusion with first-order consumption on a masked 2-D/3-D grid (sparsetype "2Dmap"/"3Dmap",
// src/testMap.f); only the cells present in the map are unknowns, the map itself arrives in ipar (model_math.h).
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_DMAP_NPAR)

__device__ void diffmap(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    const int *inv = RSM_DMAP_INV(ip), *pres = RSM_DMAP_PRES(ip);
    const int n = *neq, nthreads = (int)RS_THREAD_COUNT;
    for (int c = (int)RS_THREAD_ID; c < n; c += nthreads) ydot[c] = rsm_diffmap_point(y, c, inv, pres, d1, d2, d3, parms);
}

RS_MODEL(diffmap, RSM_DMAP_NPAR)
