/*
 * rootsolve_b200_model.cuh -- how a compiled model plugs into the GPU solver.
 *
 * rootSolve's compiled models are C/Fortran functions in a DLL with the
 * signature (src/steady.h:34, vignettes/rootSolve.Rnw:888-934)
 *
 *     void func(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
 *
 * plus an optional initialiser  void initfunc(void (*steadyparms)(int *, double *))
 * that receives the parameter vector (src/steady.h:35, inst/dynload/anoxmodc.c:21-25).
 * Here the same two functions are written once in a .cu file:
 *
 *     #include "rootsolve_b200_model.cuh"
 *     RS_MODEL_PARMS(3)                       // double parms[3] in __constant__ memory
 *     __device__ void mymod(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
 *     {
 *         RS_FOR_EACH(i, *neq) ydot[i] = -parms[0] * y[i];    // each thread does its share
 *     }
 *     RS_MODEL(mymod, 3)                      // exports  rs_model_mymod()  from the model .so
 *
 * Contract of the device function (what differs from the CPU DLL):
 *   * it is executed by EVERY thread of a (blocks x batch) grid, all with the
 *     same six arguments; it must compute ydot[i] for the indices it owns --
 *     RS_FOR_EACH / RS_THREAD_ID / RS_THREAD_COUNT partition 0..n-1 over the
 *     threads of one evaluation -- and must not write y;
 *   * y and ydot always use GLOBAL indices (the layout of the R state vector),
 *     also when the solve is slab-partitioned over several GPUs: a rank then
 *     owns planes [RS_SLAB_J0(ip), RS_SLAB_J1(ip)) of the slowest grid index
 *     and must compute ydot only there (its y holds valid values on those
 *     planes and one halo plane on either side);
 *   * ip[] and yout[] have the reference's layout (src/steady_utils.c:109-124):
 *     ip[0]=nout, ip[1]=length(yout), ip[2]=length(ip)=3+length(ipar), ip[3..]=ipar;
 *     yout[0..nout) outputs, then rpar.  The two ints after the user's ipar are
 *     the slab bounds (J1 < 0: not partitioned).
 *
 * The kernel that calls func is instantiated in the model's own translation
 * unit (template argument = the function), so the residual is inlined into
 * the batched launch; the solver library reaches it through the host function
 * pointer in rs_model_desc.  This is the GPU analogue of R's
 * getNativeSymbolInfo(func, dllname)$address (R/stodes.R:212-220).
 */
#ifndef ROOTSOLVE_B200_MODEL_CUH
#define ROOTSOLVE_B200_MODEL_CUH

#include <cuda_runtime.h>

#define RS_MODEL_ABI 1

typedef void rs_deriv_func_t(int *, double *, double *, double *, double *, int *);
typedef void rs_init_func_t(void (*)(int *, double *));

struct rs_model_desc {
    int abi;
    const char *name;
    int nparms; /* length the initialiser asks for; 0 = no initialiser */
    /* Evaluates nbatch residuals in ONE launch: evaluation b reads y + b*ld, writes ydot + b*ld and
     * uses yout + b*ldout.  work = number of unknowns this rank owns (sizes the grid). */
    int (*launch)(void *stream, int neq, double t, const double *y, double *ydot, long long ld, int nbatch,
                  double *yout, long long ldout, int *ip, long long work);
    rs_init_func_t *initfunc; /* reference signature; copies parms to the device */
    rs_init_func_t *initforc;
};

#define RS_THREAD_ID ((long long)blockIdx.x * blockDim.x + threadIdx.x)
#define RS_THREAD_COUNT ((long long)gridDim.x * blockDim.x)
#define RS_FOR_EACH(i, n) for (long long i = RS_THREAD_ID; i < (long long)(n); i += RS_THREAD_COUNT)
#define RS_SLAB_J0(ip) ((ip)[(ip)[2]])
#define RS_SLAB_J1(ip) ((ip)[(ip)[2] + 1])

template <rs_deriv_func_t F>
__global__ void __launch_bounds__(256) rs_residual_batched_kernel(int neq, double t, const double *y,
                                                                   double *ydot, long long ld, double *yout,
                                                                   long long ldout, int *ip)
{
    const long long b = blockIdx.y;
    int n = neq;
    double tt = t;
    F(&n, &tt, const_cast<double *>(y) + b * ld, ydot + b * ld, yout + b * ldout, ip);
}

template <rs_deriv_func_t F>
static int rs_launch_batched(void *stream, int neq, double t, const double *y, double *ydot, long long ld,
                             int nbatch, double *yout, long long ldout, int *ip, long long work)
{
    if (work <= 0 || nbatch <= 0) return 0;
    const int threads = 256;
    long long blocks = (work + threads - 1) / threads;
    /* One wave: the whole (blocks x nbatch) grid is resident at once -- SMs x the CTAs of THIS model's kernel an SM
     * holds (its register count decides) -- and every thread strides over its share; a grid that is a fraction of
     * a wave larger leaves the SMs idle for most of a second wave. */
    static int resident = 0;
    if (resident == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rs_residual_batched_kernel<F>, threads, 0) == cudaSuccess &&
            sms > 0 && per_sm > 0)
            resident = sms * per_sm;
        else
            resident = 148 * 4;
    }
    long long cap = resident / nbatch;
    if (cap < 1) cap = 1;
    if (blocks > cap) blocks = cap;
    dim3 grid((unsigned)blocks, (unsigned)nbatch);
    rs_residual_batched_kernel<F><<<grid, threads, 0, (cudaStream_t)stream>>>(neq, t, y, ydot, ld, yout, ldout, ip);
    return (int)cudaGetLastError();
}

/* N parameters in __constant__ memory + the reference-style initialiser that fills them */
#define RS_MODEL_PARMS(N)                                                        \
    static __constant__ double parms[N];                                         \
    static double rs_parms_host_[N];                                             \
    static void rs_model_init_(void (*steadyparms)(int *, double *))             \
    {                                                                            \
        int n_ = N;                                                              \
        steadyparms(&n_, rs_parms_host_);                                        \
        cudaMemcpyToSymbol(parms, rs_parms_host_, sizeof(double) * N);           \
    }

/* N forcing function values in __constant__ memory + the initialiser `initforc` of the reference's compiled
 * models (deSolve's convention; R/stodes.R:222-237, src/steady_utils.c:62-77).  A steady-state solve sees the
 * forcings at ONE time, so the values are constants of the solve. */
#define RS_MODEL_FORCS(N)                                                        \
    static __constant__ double forcs[N];                                         \
    static double rs_forcs_host_[N];                                             \
    static void rs_model_initforc_(void (*steadyforcs)(int *, double *))         \
    {                                                                            \
        int n_ = N;                                                              \
        steadyforcs(&n_, rs_forcs_host_);                                        \
        cudaMemcpyToSymbol(forcs, rs_forcs_host_, sizeof(double) * N);           \
    }

#define RS_MODEL_NAMED_F(NAME, FUNC, NPARMS, INITFN, FORCFN)                     \
    extern "C" const rs_model_desc *rs_model_##NAME(void)                        \
    {                                                                            \
        static const rs_model_desc d = {RS_MODEL_ABI, #NAME, NPARMS, rs_launch_batched<FUNC>, INITFN, FORCFN}; \
        return &d;                                                               \
    }
#define RS_MODEL_NAMED(NAME, FUNC, NPARMS, INITFN) RS_MODEL_NAMED_F(NAME, FUNC, NPARMS, INITFN, nullptr)

/* model with RS_MODEL_PARMS(NPARMS) declared earlier in the same file */
#define RS_MODEL(FUNC, NPARMS) RS_MODEL_NAMED(FUNC, FUNC, NPARMS, rs_model_init_)
/* model with RS_MODEL_PARMS(NPARMS) and RS_MODEL_FORCS(...) declared earlier in the same file */
#define RS_MODEL_WITH_FORCS(FUNC, NPARMS) RS_MODEL_NAMED_F(FUNC, FUNC, NPARMS, rs_model_init_, rs_model_initforc_)
/* model without parameters */
#define RS_MODEL_NOPARMS(FUNC) RS_MODEL_NAMED(FUNC, FUNC, 0, nullptr)

#endif /* ROOTSOLVE_B200_MODEL_CUH */
