// Bundled model: diff3d plus the source term of a manufactured solution (closed-form discrete steady state at any
// grid size; SURVEY.md 8(c), parity protocol step 3).  The three sine tables arrive in rpar.
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_D3MS_NPAR)

__device__ void diff3dms(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    const double *tab = yout + ip[0];
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const rsm_d3_coef c = rsm_diff3d_coef(parms);
    const double c0 = parms[8], c1 = parms[9];
    RS_FOR_EACH_CELL(it, l, j0, j1, d2, d3)
        ydot[(it.j * d2 + it.k) * d3 + l] = rsm_diff3dms_point(y, it.j, it.k, l, d1, d2, d3, c, c0, c1, tab);
}

__device__ double diff3dms_row(long long i, double t, rs_state y, const double *yout, const int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    const int plane = d2 * d3, idx = (int)i;
    const int j = idx / plane, rem = idx - j * plane, k = rem / d3, l = rem - k * d3;
    return rsm_diff3dms_point(y, j, k, l, d1, d2, d3, rsm_diff3d_coef(parms), parms[8], parms[9], yout + ip[0]);
}

RS_MODEL_ACC_ROW(diff3dms, diff3dms_row, RSM_D3MS_NPAR)
