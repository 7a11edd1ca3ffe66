// Bundled model: multi-species 1-D reaction-transport column (BASELINE config C2).
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_RT1_NPAR)

__device__ void rt1d(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int nspec = ip[3], nb = ip[4];
    int j0 = 0, j1 = nb;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const long long count = j1 - j0;
    RS_FOR_EACH(q, count) {
        const int j = j0 + (int)q;
        for (int s = 0; s < nspec; s++) ydot[(long long)s * nb + j] = rsm_rt1d_point(y, s, j, nspec, nb, parms);
    }
}

RS_MODEL_ACC(rt1d, RSM_RT1_NPAR)
