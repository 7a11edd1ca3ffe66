// example_model.cu -- a user's own model, outside the bundled library: what a rootSolve user writes instead of a
// C/Fortran DLL (vignettes/rootSolve.Rnw:888-934, inst/dynload/anoxmodc.c).  Logistic growth with diffusion on a
// 1-D transect (Fisher's equation): an imposed density at the left end, zero gradient at the right end.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -shared -Xcompiler -fPIC -Iinclude \
//        integration/example_model.cu -o libfisher.so
//
//   out = rootsolve_b200.steady_1D(y0, func="fisher1d", dllname="./libfisher.so", parms=[D, dx, r, K, yleft],
//                                  nspec=1, positive=True)
//
// The function has the compiled-model signature of the reference; every thread of the launch runs it and computes
// its share of ydot (RS_FOR_EACH).  A model that should also run slab-partitioned over several GPUs restricts itself
// to the boxes [RS_SLAB_J0, RS_SLAB_J1) it is given, as below.
#include "rootsolve_b200_model.cuh"

RS_MODEL_PARMS(5)   // D dx r K yleft -- filled by the initialiser, as `parms' of an R call reaches a DLL's initfunc

__device__ void fisher1d(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    const int n = *neq;
    const double D = parms[0], dx = parms[1], r = parms[2], K = parms[3], yleft = parms[4];
    int j0 = 0, j1 = n;
    if (RS_SLAB_J1(ip) >= 0) { j0 = RS_SLAB_J0(ip); j1 = RS_SLAB_J1(ip); }
    RS_FOR_EACH(q, j1 - j0) {
        const int i = j0 + (int)q;
        const double yc = y[i];
        const double lo = (i > 0) ? y[i - 1] : yleft;
        const double hi = (i < n - 1) ? y[i + 1] : yc;
        ydot[i] = D * (hi - 2.0 * yc + lo) / (dx * dx) + r * yc * (1.0 - yc / K);
    }
}

RS_MODEL(fisher1d, 5)
