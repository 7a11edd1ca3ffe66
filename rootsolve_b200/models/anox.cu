// Bundled model: the 4-variable anoxic sediment model the reference ships as its compiled-code example
// (inst/dynload/anoxmodc.c, vignettes/rootSolve.Rnw:352-372), with its one output variable SumS.
// A tiny system: one thread of each evaluation does all of it.
#include "rootsolve_b200_model.cuh"
#include "model_math.h"

RS_MODEL_PARMS(RSM_ANOX_NPAR)

__device__ void anoxmod(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    if (RS_THREAD_ID != 0) return;
    if (ip[0] < 1) return;   // "nout should be at least 1" (inst/dynload/anoxmodc.c:36): checked on the host side
    rsm_anox(y, ydot, yout, parms);
}

RS_MODEL(anoxmod, RSM_ANOX_NPAR)
