/*
 * oracle_models.c -- host-side compiled models with the rootSolve DLL signature
 * func(neq, t, y, ydot, yout, ip) (src/steady.h:34, inst/dynload/anoxmodc.c:29-53)
 * for the oracle and the CPU baseline.  TEST INFRASTRUCTURE ONLY.
 *
 * The point-wise math is shared with the CUDA models through
 * rootsolve_b200/models/model_math.h (see the note there); these wrappers are
 * the plain sequential "whole vector per call" form the reference expects.
 * Parameters arrive the way compiled rootSolve models get them: an initialiser
 * init<model>(steadyparms) that calls back steadyparms(&N, parms)
 * (src/steady_utils.c:43-58, inst/dynload/anoxmodc.c:21-25).
 */
#include "../rootsolve_b200/models/model_math.h"

static double p_d2[RSM_D2_NPAR], p_d3[RSM_D3_NPAR], p_rd3[RSM_RD3_NPAR], p_rt1[RSM_RT1_NPAR];

void initdiff2d(void (*steadyparms)(int *, double *)) { int N = RSM_D2_NPAR; steadyparms(&N, p_d2); }
void initdiff3d(void (*steadyparms)(int *, double *)) { int N = RSM_D3_NPAR; steadyparms(&N, p_d3); }
void initrd3d(void (*steadyparms)(int *, double *)) { int N = RSM_RD3_NPAR; steadyparms(&N, p_rd3); }
void initrt1d(void (*steadyparms)(int *, double *)) { int N = RSM_RT1_NPAR; steadyparms(&N, p_rt1); }

void diff2d(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)yout;
    const int d1 = ip[4], d2 = ip[5], variant = ip[7];
    const rsm_d2_coef c = rsm_diff2d_coef(p_d2);
    for (int j = 0; j < d1; j++)
        for (int k = 0; k < d2; k++) ydot[j * d2 + k] = rsm_diff2d_point(y, j, k, d1, d2, c, variant);
}

void diff3d(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)yout;
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6], variant = ip[7];
    const rsm_d3_coef c = rsm_diff3d_coef(p_d3);
    for (int j = 0; j < d1; j++)
        for (int k = 0; k < d2; k++)
            for (int l = 0; l < d3; l++)
                ydot[(j * d2 + k) * d3 + l] = rsm_diff3d_point(y, j, k, l, d1, d2, d3, c, variant);
}

void rd3d(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)yout;
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    const long nc = (long)d1 * d2 * d3;
    const rsm_rd3_coef c = rsm_rd3d_coef(p_rd3);
    for (int s = 0; s < 2; s++)
        for (int j = 0; j < d1; j++)
            for (int k = 0; k < d2; k++)
                for (int l = 0; l < d3; l++)
                    ydot[s * nc + (j * d2 + k) * d3 + l] = rsm_rd3d_point(y, s, j, k, l, d1, d2, d3, c);
}

void rt1d(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)yout;
    const int nspec = ip[3], nb = ip[4];
    for (int s = 0; s < nspec; s++)
        for (int j = 0; j < nb; j++) ydot[(long)s * nb + j] = rsm_rt1d_point(y, s, j, nspec, nb, p_rt1);
}

/* the reference's compiled-code example, inst/dynload/anoxmodc.c (4 state variables, 1 output) */
static double p_anox[RSM_ANOX_NPAR];
void initanoxmod(void (*steadyparms)(int *, double *)) { int N = RSM_ANOX_NPAR; steadyparms(&N, p_anox); }
void anoxmod(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)ip;
    rsm_anox(y, ydot, yout, p_anox);
}

/* ... and with Flux and BO2 as forcing function values (initforc convention, src/steady_utils.c:62-77) */
static double p_anoxf[RSM_ANOX_NPAR], f_anoxf[RSM_ANOXF_NFORC];
void initanoxforc(void (*steadyparms)(int *, double *)) { int N = RSM_ANOX_NPAR; steadyparms(&N, p_anoxf); }
void forcanoxforc(void (*steadyforcs)(int *, double *)) { int N = RSM_ANOXF_NFORC; steadyforcs(&N, f_anoxf); }
void anoxforc(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)ip;
    double p[RSM_ANOX_NPAR];
    rsm_anox_forced_parms(p_anoxf, f_anoxf, p);
    rsm_anox(y, ydot, yout, p);
}

/* man/steady.2D.Rd:167-206, the cyclic example */
static double p_d2b[RSM_D2B_NPAR];
void initdiff2db(void (*steadyparms)(int *, double *)) { int N = RSM_D2B_NPAR; steadyparms(&N, p_d2b); }
void diff2db(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)yout;
    const int d1 = ip[4], d2 = ip[5];
    for (int j = 0; j < d1; j++)
        for (int k = 0; k < d2; k++) ydot[j * d2 + k] = rsm_diff2db_point(y, j, k, d1, d2, p_d2b);
}

/* man/steady.1D.Rd:113-153, BOD and oxygen in a river; yout[0..nout) = the BOD decay rate per box */
static double p_o2bod[RSM_O2BOD_NPAR];
void inito2bod(void (*steadyparms)(int *, double *)) { int N = RSM_O2BOD_NPAR; steadyparms(&N, p_o2bod); }
void o2bod(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t;
    const int nb = ip[4], nout = ip[0];
    for (int s = 0; s < 2; s++)
        for (int j = 0; j < nb; j++) ydot[s * nb + j] = rsm_o2bod_point(y, s, j, nb, p_o2bod);
    for (int j = 0; j < nb && j < nout; j++) yout[j] = rsm_o2bod_rate(y, j, nb, p_o2bod);
}

/* diffusion on a masked grid (sparsetype "2Dmap"/"3Dmap"); the map is in ipar, see model_math.h */
static double p_dmap[RSM_DMAP_NPAR];
void initdiffmap(void (*steadyparms)(int *, double *)) { int N = RSM_DMAP_NPAR; steadyparms(&N, p_dmap); }
void diffmap(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)t; (void)yout;
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    for (int c = 0; c < *neq; c++)
        ydot[c] = rsm_diffmap_point(y, c, RSM_DMAP_INV(ip), RSM_DMAP_PRES(ip), d1, d2, d3, p_dmap);
}

/* manufactured-solution variant of diff3d: sine tables in rpar (= yout + nout) */
static double p_d3ms[RSM_D3MS_NPAR];
void initdiff3dms(void (*steadyparms)(int *, double *)) { int N = RSM_D3MS_NPAR; steadyparms(&N, p_d3ms); }
void diff3dms(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t;
    const int d1 = ip[4], d2 = ip[5], d3 = ip[6];
    const double *tab = yout + ip[0];
    const rsm_d3_coef c = rsm_diff3d_coef(p_d3ms);
    for (int j = 0; j < d1; j++)
        for (int k = 0; k < d2; k++)
            for (int l = 0; l < d3; l++)
                ydot[(j * d2 + k) * d3 + l] = rsm_diff3dms_point(y, j, k, l, d1, d2, d3, c, p_d3ms[8], p_d3ms[9], tab);
}

/* vignettes/rootSolve.Rnw:615-663, zero-flux 2-D diffusion with quadratic consumption and point sources */
static double p_d2q[RSM_D2Q_NPAR];
void initdiff2dq(void (*steadyparms)(int *, double *)) { int N = RSM_D2Q_NPAR; steadyparms(&N, p_d2q); }
void diff2dq(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)yout;
    const int d1 = ip[4], d2 = ip[5], nsrc = ip[7];
    for (int j = 0; j < d1; j++)
        for (int k = 0; k < d2; k++) ydot[j * d2 + k] = rsm_diff2dq_point(y, j, k, d1, d2, p_d2q, ip + 8, nsrc);
}

/* man/stodes.Rd:225-252, the 1000-layer sediment model with non-local exchange */
static double p_sed[RSM_SED_NPAR];
void initsediment(void (*steadyparms)(int *, double *)) { int N = RSM_SED_NPAR; steadyparms(&N, p_sed); }
void sediment(int *neq, double *t, double *y, double *ydot, double *yout, int *ip)
{
    (void)neq; (void)t; (void)yout;
    const int N = ip[4];
    for (int i = 0; i < N; i++) ydot[i] = rsm_sediment_point(y, i, N, p_sed);
}
