// Bundled model: the anoxic sediment model (anox.cu) with two of its constants supplied as FORCING FUNCTIONS the
// way deSolve-style compiled models take them (R/stodes.R:222-237, src/steady_utils.c:62-77): forcs[0] replaces
// the organic-matter flux, forcs[1] the bottom-water oxygen concentration.  The slots parms[1] and parms[6] are
// ignored.  One output variable, SumS.
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_ANOX_NPAR)
RS_MODEL_FORCS(RSM_ANOXF_NFORC)

__device__ void anoxforc(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    if (RS_THREAD_ID != 0) return;
    if (ip[0] < 1) return;
    double p[RSM_ANOX_NPAR];
    rsm_anox_forced_parms(parms, forcs, p);
    rsm_anox(y, ydot, yout, p);
}

RS_MODEL_WITH_FORCS(anoxforc, RSM_ANOX_NPAR)
