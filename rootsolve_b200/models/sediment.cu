// Bundled model: the 1000-equation sediment example of man/stodes.Rd:225-252 (advection + bioturbation + decay,
// non-local exchange from the top five layers to layer N/2).
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_SED_NPAR)

__device__ void sediment(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int N = ip[4];
    RS_FOR_EACH(i, N) ydot[i] = rsm_sediment_point(y, (int)i, N, parms);
}

RS_MODEL_ACC(sediment, RSM_SED_NPAR)
