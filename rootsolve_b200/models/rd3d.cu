// Bundled model: two-species reaction-diffusion on a 3-D grid (BASELINE config C4).
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_RD3_NPAR)

__device__ void rd3d(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    int j0 = 0, j1 = d1;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    const int nc = d1 * d2 * d3;
    const rsm_rd3_coef c = rsm_rd3d_coef(parms);
    RS_FOR_EACH_CELL(it, l, j0, j1, d2, d3) {
        const int cell = (it.j * d2 + it.k) * d3 + l;
        ydot[cell] = rsm_rd3d_point(y, 0, it.j, it.k, l, d1, d2, d3, c);
        ydot[nc + cell] = rsm_rd3d_point(y, 1, it.j, it.k, l, d1, d2, d3, c);
    }
}

__device__ double rd3d_row(long long i, double t, rs_state y, const double *yout, const int *ip)
{
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    const int plane = d2 * d3, nc = plane * d1, idx = (int)i;
    const int s = idx >= nc ? 1 : 0, cell = idx - s * nc;
    const int j = cell / plane, rem = cell - j * plane, k = rem / d3, l = rem - k * d3;
    return rsm_rd3d_point(y, s, j, k, l, d1, d2, d3, rsm_rd3d_coef(parms));
}

RS_MODEL_ACC_ROW(rd3d, rd3d_row, RSM_RD3_NPAR)
