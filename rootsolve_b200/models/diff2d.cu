// Bundled model: 2-D diffusion with first-order or Monod consumption and imposed
// boundary values (man/steady.2D.Rd:110-125; BASELINE configs C1 and C3).
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_D2_NPAR)

__device__ void diff2d(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], variant = ip[7];
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const rsm_d2_coef c = rsm_diff2d_coef(parms);
    RS_FOR_EACH_CELL(it, k, j0, j1, 1, d2)
        ydot[it.j * d2 + k] = rsm_diff2d_point(y, it.j, k, d1, d2, c, variant);
}

__device__ double diff2d_row(long long i, double t, rs_state y, const double *yout, const int *ip)
{
    const int d1 = ip[4], d2 = ip[5], variant = ip[7], idx = (int)i;
    const int j = idx / d2, k = idx - j * d2;
    return rsm_diff2d_point(y, j, k, d1, d2, rsm_diff2d_coef(parms), variant);
}

RS_MODEL_ACC_ROW(diff2d, diff2d_row, RSM_D2_NPAR)
