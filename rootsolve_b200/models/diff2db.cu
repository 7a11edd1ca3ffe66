// Bundled model: the second example of man/steady.2D.Rd (`diffusion2Db`, :167-206): 2-D diffusion with consumption,
// an extra flux on two sides and a one-way exchange across the boundary the example declares cyclic (cyclicBnd = 2).
// ipar as the other grid models: ip[3]=nspec ip[4]=d1 ip[5]=d2.
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_D2B_NPAR)

__device__ void diff2db(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5];
    RS_FOR_EACH_CELL(it, k, 0, d1, 1, d2)
        ydot[it.j * d2 + k] = rsm_diff2db_point(y, it.j, k, d1, d2, parms);
}

RS_MODEL_ACC(diff2db, RSM_D2B_NPAR)
