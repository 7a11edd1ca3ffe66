// Bundled model: 3-D diffusion with first-order consumption and imposed boundary
// values (man/steady.3D.Rd:105-140, vignettes/rootSolve.Rnw:686-739; BASELINE
// config C5 and the 10^3 golden case).  Accessor form: the perturbed states of the
// finite-difference Jacobian are never materialised.
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_D3_NPAR)

__device__ void diff3d(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6], variant = ip[7];
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const rsm_d3_coef c = rsm_diff3d_coef(parms);
    RS_FOR_EACH_CELL(it, l, j0, j1, d2, d3)
        ydot[(it.j * d2 + it.k) * d3 + l] = rsm_diff3d_point(y, it.j, it.k, l, d1, d2, d3, c, variant);
}

// one row of the same function: what the fused finite-difference Jacobian evaluates (row form of the SDK)
__device__ double diff3d_row(long long i, double t, rs_state y, const double *yout, const int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6], variant = ip[7];
    const int plane = d2 * d3, idx = (int)i;
    const int j = idx / plane, rem = idx - j * plane, k = rem / d3, l = rem - k * d3;
    return rsm_diff3d_point(y, j, k, l, d1, d2, d3, rsm_diff3d_coef(parms), variant);
}

RS_MODEL_ACC_ROW(diff3d, diff3d_row, RSM_D3_NPAR)
