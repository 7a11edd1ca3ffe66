// Bundled model: the first example of man/steady.1D.Rd (`O2BOD`, :113-153) -- BOD and oxygen along a river, two
// species, upwind transport, Monod-limited decay -- with its output vector BODrate (nout = number of boxes).
// ipar: ip[3]=nspec (2) ip[4]=number of boxes.  Sharded solves: a rank computes the boxes of its slab only.
#include "acc_prelude.h"

RS_MODEL_PARMS(RSM_O2BOD_NPAR)

__device__ void o2bod(int *neq, double *t, rs_state y, double *ydot, double *yout, int *ip)
{
    const int nb = ip[4], nout = ip[0];
    int j0 = 0, j1 = nb;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    RS_FOR_EACH(q, j1 - j0) {
        const int j = j0 + (int)q;
        ydot[j] = rsm_o2bod_point(y, 0, j, nb, parms);
        ydot[nb + j] = rsm_o2bod_point(y, 1, j, nb, parms);
        if (j < nout) yout[j] = rsm_o2bod_rate(y, j, nb, parms);
    }
}

RS_MODEL_ACC(o2bod, RSM_O2BOD_NPAR)
