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

RS_MODEL_ACC(diff3d, RSM_D3_NPAR)
