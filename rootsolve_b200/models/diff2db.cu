// Bundled model: the second example of man/steady.2D.Rd (`diffusion2Db`, :167-206): 2-D diffusion with consumption,
// an extra flux on two sides and a one-way exchange across the boundary the example declares cyclic (cyclicBnd = 2).
// ipar as the other grid models: ip[3]=nspec ip[4]=d1 ip[5]=d2.
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_D2B_NPAR)

__device__ void diff2db(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5];
    const int n = d1 * d2, nthreads = (int)RS_THREAD_COUNT;
    for (int idx = (int)RS_THREAD_ID; idx < n; idx += nthreads) {
        const int j = idx / d2, k = idx - j * d2;
        ydot[idx] = rsm_diff2db_point(y, j, k, d1, d2, parms);
    }
}

RS_MODEL(diff2db, RSM_D2B_NPAR)
