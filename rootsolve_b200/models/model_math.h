/*
 * model_math.h -- point-wise right-hand sides of the bundled synthetic
 * reaction-transport models (BASELINE.json configs C1..C5).
 *
 * One source, two compilers: nvcc builds these as __device__ functions for the
 * CUDA models (bundled_models.cu); gcc builds them as plain C for the host
 * wrappers the oracle and the CPU baseline call (oracle/oracle_models.c).  The
 * models are *user input* to the solver (what rootSolve users write as an R
 * closure or a compiled DLL func), not part of the reference's algorithm, so
 * sharing the source between checker and product is deliberate: it makes the
 * residual bit-identical on both sides (device build uses -fmad=false) and
 * leaves the solver itself as the only thing the parity tests compare.
 *
 * Layout (what rootSolve's R wrappers pass down, R/steady.2D.R:48-49,
 * R/steady.3D.R:49-50, src/sparsecoeff.f:96,125,215,244): species-major; within
 * a species the grid is stored with the *last* stodes dimension fastest,
 *   2-D: idx = s*d1*d2 + j*d2 + k            (j in [0,d1), k in [0,d2))
 *   3-D: idx = s*d1*d2*d3 + j*d2*d3 + k*d3 + l
 * where (d1,d2[,d3]) = rev(dimens) of the R call.
 *
 * ipar convention of the bundled models (after the 3 reserved slots of
 * src/steady_utils.c:109-124): ip[3]=nspec, ip[4]=d1, ip[5]=d2, ip[6]=d3,
 * ip[7]=variant flag.
 */
#ifndef RSM_MODEL_MATH_H
#define RSM_MODEL_MATH_H

#ifndef RSM_HD
#ifdef __CUDACC__
#define RSM_HD __host__ __device__ __forceinline__
#else
#define RSM_HD static inline
#endif
#endif

/* The state argument of the point functions: a plain vector for the host wrappers; the device models define it as
 * the SDK's rs_state before including this file, so that y[i] carries the finite-difference increment of the
 * evaluation on the fly (rootsolve_b200_model.cuh, accessor form).  Both support y[i] and y + offset. */
#ifndef RSM_STATE
#define RSM_STATE const double *
#endif

/* ---- parameter slots ------------------------------------------------- */
/* diff2d (C1/C3 and man/steady.2D.Rd:110-125): 7 parameters */
#define RSM_D2_NPAR 7
/* p[0]=Dx p[1]=Dy p[2]=dx p[3]=dy p[4]=r (linear rate, or rmax) p[5]=ks p[6]=BND */
/* diff3d (C5 and man/steady.3D.Rd:105-140): 8 parameters */
#define RSM_D3_NPAR 8
/* p[0]=Dx p[1]=Dy p[2]=Dz p[3]=dx p[4]=dy p[5]=dz p[6]=r p[7]=BND */
/* rd3d two-species (C4): 10 parameters */
#define RSM_RD3_NPAR 10
/* p[0]=D1 p[1]=D2 p[2]=dx p[3]=dy p[4]=dz p[5]=k p[6]=m p[7]=BNDu p[8]=BNDv p[9]=unused */
/* rt1d multi-species reaction-transport (C2): 6 parameters */
#define RSM_RT1_NPAR 7
/* p[0]=v p[1]=D p[2]=dx p[3]=r p[4]=ks p[5]=BND (upstream, all species) p[6]=r0 */

/* One direction of Fickian diffusion with imposed (Dirichlet) values outside
 * the grid, written in the flux form the man-page models use:
 *   Flux_lo = -D*(yc - ylo)/h ; Flux_hi = -D*(yhi - yc)/h ; -(Flux_hi-Flux_lo)/h
 * with ylo/yhi replaced by BND at the edges. */
RSM_HD double rsm_flux_div(double ylo, double yc, double yhi, double D, double h)
{
    double flo = -D * (yc - ylo) / h;
    double fhi = -D * (yhi - yc) / h;
    return (fhi - flo) / h;
}

/* The same flux divergence with the coefficient w = D / h^2 formed once per model call instead of three divisions
 * per direction and cell: (Flux_hi - Flux_lo)/h = w * ((yc - ylo) - (yhi - yc)).  For D = h = 1 (the reference's
 * examples, the goldens) both forms round identically. */
RSM_HD double rsm_flux_div_w(double ylo, double yc, double yhi, double w)
{
    return w * ((yc - ylo) - (yhi - yc));
}

/* variant: 0 = first-order consumption -r*y ; 1 = Monod -rmax*y/(y+ks) */
typedef struct { double wx, wy, r, ks, bnd; } rsm_d2_coef;
RSM_HD rsm_d2_coef rsm_diff2d_coef(const double *p)
{
    rsm_d2_coef c;
    c.wx = p[0] / (p[2] * p[2]); c.wy = p[1] / (p[3] * p[3]); c.r = p[4]; c.ks = p[5]; c.bnd = p[6];
    return c;
}
RSM_HD double rsm_diff2d_point(RSM_STATE y, int j, int k, int d1, int d2, const rsm_d2_coef c, int variant)
{
    const int idx = j * d2 + k;
    const double yc = y[idx];
    double dY = (variant == 1) ? -c.r * yc / (yc + c.ks) : -c.r * yc;
    /* "X" = first R dimension = fastest index k */
    double ylo = (k > 0) ? y[idx - 1] : c.bnd;
    double yhi = (k < d2 - 1) ? y[idx + 1] : c.bnd;
    dY = dY - rsm_flux_div_w(ylo, yc, yhi, c.wx);
    /* "Y" = second R dimension = j */
    ylo = (j > 0) ? y[idx - d2] : c.bnd;
    yhi = (j < d1 - 1) ? y[idx + d2] : c.bnd;
    dY = dY - rsm_flux_div_w(ylo, yc, yhi, c.wy);
    return dY;
}

/* The second example of man/steady.2D.Rd (:167-206, `diffusion2Db`): the same diffusion-consumption model with an
 * extra flux on two sides and an exchange between the two sides of the second R dimension -- the one the example
 * declares cyclic (cyclicBnd = 2).  The exchange is one-way (the last column of the R matrix is driven by the first),
 * so the Jacobian is nonsymmetric and uses one of the two wrap entries of the cyclic pattern.
 * p[0]=Dx p[1]=Dy p[2]=dx p[3]=dy p[4]=r p[5]=BND p[6]=extra flux (10) p[7]=exchange rate (10)
 * j = index of the second R dimension (slow), k = index of the first (fast), as everywhere in this file. */
#define RSM_D2B_NPAR 8
RSM_HD double rsm_diff2db_point(RSM_STATE y, int j, int k, int d1, int d2, const double *p)
{
    const double Dx = p[0], Dy = p[1], dx = p[2], dy = p[3], r = p[4], bnd = p[5], extra = p[6], exch = p[7];
    const int idx = j * d2 + k;
    const double yc = y[idx];
    double dY = -r * yc;
    dY = dY - rsm_flux_div((k > 0) ? y[idx - 1] : bnd, yc, (k < d2 - 1) ? y[idx + 1] : bnd, Dx, dx);
    dY = dY - rsm_flux_div((j > 0) ? y[idx - d2] : bnd, yc, (j < d1 - 1) ? y[idx + d2] : bnd, Dy, dy);
    if (j == 0) dY = dY + extra;                              /* dY[,1]  <- dY[,1] + 10 */
    if (k == 0) dY = dY + extra;                              /* dY[1,]  <- dY[1,] + 10 */
    if (j == d1 - 1) dY = dY + (y[k] - yc) * exch;            /* dY[,ny] <- dY[,ny] + (y[,1] - y[,ny]) * 10 */
    return dY;
}

/* variant: 0 = each direction once ; 1 = each direction applied twice, in the
 * order the vignette's three loops apply them (vignettes/rootSolve.Rnw:686-722):
 * (y,z) for fixed x-slices, (x,z), (x,y). */
typedef struct { double wx, wy, wz, r, bnd; } rsm_d3_coef;
RSM_HD rsm_d3_coef rsm_diff3d_coef(const double *p)
{
    rsm_d3_coef c;
    c.wx = p[0] / (p[3] * p[3]); c.wy = p[1] / (p[4] * p[4]); c.wz = p[2] / (p[5] * p[5]); c.r = p[6]; c.bnd = p[7];
    return c;
}
RSM_HD double rsm_diff3d_point(RSM_STATE y, int j, int k, int l, int d1, int d2, int d3, const rsm_d3_coef c, int variant)
{
    const double r = c.r, bnd = c.bnd;
    const int s2 = d3, s1 = d2 * d3;
    const int idx = j * s1 + k * s2 + l;
    const double yc = y[idx];
    double dY = -r * yc;
    /* x = first R dim = l ; y = second R dim = k ; z = third R dim = j */
    const double gx = rsm_flux_div_w((l > 0) ? y[idx - 1] : bnd, yc, (l < d3 - 1) ? y[idx + 1] : bnd, c.wx);
    const double gy = rsm_flux_div_w((k > 0) ? y[idx - s2] : bnd, yc, (k < d2 - 1) ? y[idx + s2] : bnd, c.wy);
    const double gz = rsm_flux_div_w((j > 0) ? y[idx - s1] : bnd, yc, (j < d1 - 1) ? y[idx + s1] : bnd, c.wz);
    if (variant == 1) {
        dY = dY - gy; dY = dY - gz;
        dY = dY - gx; dY = dY - gz;
        dY = dY - gx; dY = dY - gy;
    } else {
        dY = dY - gx; dY = dY - gy; dY = dY - gz;
    }
    return dY;
}

/* Two species u (s=0), v (s=1) on a 3-D grid:
 *   du = D1 lap(u) - k u v ;  dv = D2 lap(v) + k u v - m v ; Dirichlet faces. */
typedef struct { double wx[2], wy[2], wz[2], kk, m, bnd[2]; } rsm_rd3_coef;
RSM_HD rsm_rd3_coef rsm_rd3d_coef(const double *p)
{
    rsm_rd3_coef c;
    for (int s = 0; s < 2; s++) {
        c.wx[s] = p[s] / (p[2] * p[2]); c.wy[s] = p[s] / (p[3] * p[3]); c.wz[s] = p[s] / (p[4] * p[4]);
        c.bnd[s] = p[7 + s];
    }
    c.kk = p[5]; c.m = p[6];
    return c;
}
RSM_HD double rsm_rd3d_point(RSM_STATE y, int s, int j, int k, int l, int d1, int d2, int d3, const rsm_rd3_coef c)
{
    const double bnd = c.bnd[s];
    const int s2 = d3, s1 = d2 * d3, nc = d1 * d2 * d3;
    const int cell = j * s1 + k * s2 + l;
    RSM_STATE ys = y + (long)s * nc;
    const double yc = ys[cell];
    const double u = y[cell], v = y[nc + cell];
    const double react = c.kk * u * v;
    double dY = (s == 0) ? -react : (react - c.m * v);
    dY = dY - rsm_flux_div_w((l > 0) ? ys[cell - 1] : bnd, yc, (l < d3 - 1) ? ys[cell + 1] : bnd, c.wx[s]);
    dY = dY - rsm_flux_div_w((k > 0) ? ys[cell - s2] : bnd, yc, (k < d2 - 1) ? ys[cell + s2] : bnd, c.wy[s]);
    dY = dY - rsm_flux_div_w((j > 0) ? ys[cell - s1] : bnd, yc, (j < d1 - 1) ? ys[cell + s1] : bnd, c.wz[s]);
    return dY;
}

/* nspec species on a 1-D column of nb boxes: first-order upwind advection v,
 * diffusion D, first-order loss r0, and a Monod chain in every box in which the
 * consumption of species s is limited by the availability of species s+1,
 *   R_s = r y_s y_{s+1}/(y_{s+1}+ks)  (s < nspec-1),  R_last = r y_last,
 *   dy_s = transport - R_s - r0 y_s.
 * Upstream (box 0): imposed value BND for every species; downstream: zero
 * gradient.  The coupling is one-sided (species s depends on s+1), so the
 * Jacobian is nonsymmetric, while sparse1d still allocates the full
 * nspec x nspec cell block (src/sparsecoeff.f:70-75). */
RSM_HD double rsm_rt1d_point(RSM_STATE y, int s, int j, int nspec, int nb, const double *p)
{
    const double v = p[0], D = p[1], dx = p[2], r = p[3], ks = p[4], bnd = p[5], r0 = p[6];
    RSM_STATE ys = y + (long)s * nb;
    const double yc = ys[j];
    const double ylo = (j > 0) ? ys[j - 1] : bnd;
    const double yhi = (j < nb - 1) ? ys[j + 1] : yc; /* zero gradient */
    /* flux through the lower and upper interface of box j */
    const double flo = v * ylo - D * (yc - ylo) / dx;
    const double fhi = v * yc - D * (yhi - yc) / dx;
    double dY = -(fhi - flo) / dx;
    double Rs;
    if (s < nspec - 1) {
        const double yn = y[(long)(s + 1) * nb + j];
        Rs = r * yc * yn / (yn + ks);
    } else {
        Rs = r * yc;
    }
    dY = dY - Rs;
    dY = dY - r0 * yc;
    return dY;
}

/* The first example of man/steady.1D.Rd (:113-153, `O2BOD`): biochemical oxygen demand and oxygen in a river,
 * two species on nb boxes, upwind transport with velocity v, BOD decay limited by oxygen (Monod), re-aeration.
 * One output VECTOR: the decay rate in every box (nout = nb).
 * p[0]=v p[1]=dx p[2]=r p[3]=p p[4]=O2sat p[5]=O2_0 p[6]=BOD_0 p[7]=ksO2 (10 in the man page) */
#define RSM_O2BOD_NPAR 8
RSM_HD double rsm_o2bod_rate(RSM_STATE y, int j, int nb, const double *p)
{
    const double BOD = y[j], O2 = y[nb + j];
    return p[2] * BOD * O2 / (O2 + p[7]);
}
RSM_HD double rsm_o2bod_point(RSM_STATE y, int s, int j, int nb, const double *p)
{
    const double v = p[0], dx = p[1], rea = p[3], O2sat = p[4];
    RSM_STATE ys = y + (long)s * nb;
    const double upstream = (j > 0) ? ys[j - 1] : ((s == 0) ? p[6] : p[5]);
    const double flo = v * upstream, fhi = v * ys[j];      /* FluxX <- v * c(X_0, X) */
    const double rate = rsm_o2bod_rate(y, j, nb, p);
    double d = -(fhi - flo) / dx - rate;                   /* -diff(FluxX)/dx - BODrate */
    if (s == 1) d = d + rea * (O2sat - ys[j]);             /* + p*(O2sat-O2) */
    return d;
}

/* The 4-variable sediment model of the reference's vignette and of inst/dynload (organic matter OM, O2, SO4,
 * HS; vignettes/rootSolve.Rnw:352-372): oxic and anoxic mineralisation with Monod limitation, re-oxidation of
 * sulphide, exchange with the bottom water.  One output variable, SumS = SO4 + HS.
 * p[0]=D p[1]=Flux p[2]=r p[3]=rox p[4]=ks p[5]=ks2 p[6]=BO2 p[7]=BSO4 p[8]=BHS */
#define RSM_ANOX_NPAR 9
RSM_HD void rsm_anox(const double *y, double *ydot, double *yout, const double *p)
{
    const double D = p[0], Flux = p[1], r = p[2], rox = p[3], ks = p[4], ks2 = p[5];
    const double BO2 = p[6], BSO4 = p[7], BHS = p[8];
    const double OM = y[0], O2 = y[1], SO4 = y[2], HS = y[3];
    const double lim = O2 / (O2 + ks);
    const double oxicmin = r * OM * lim;
    const double anoxicmin = r * OM * (1.0 - lim) * SO4 / (SO4 + ks2);
    const double reox = rox * HS * lim;
    ydot[0] = Flux - oxicmin - anoxicmin;
    ydot[1] = -oxicmin - 2.0 * reox + D * (BO2 - O2);
    ydot[2] = -0.5 * anoxicmin + reox + D * (BSO4 - SO4);
    ydot[3] = 0.5 * anoxicmin - reox + D * (BHS - HS);
    yout[0] = SO4 + HS;
}

/* the same model with Flux and BO2 taken from two forcing function values (anoxforc) */
#define RSM_ANOXF_NFORC 2
RSM_HD void rsm_anox_forced_parms(const double *p, const double *forcs, double *q)
{
    for (int i = 0; i < RSM_ANOX_NPAR; i++) q[i] = p[i];
    q[1] = forcs[0];
    q[6] = forcs[1];
}

/* Diffusion with first-order consumption on a MASKED 2-D or 3-D grid (sparsetype "2Dmap"/"3Dmap", src/testMap.f):
 * the state vector holds only the cells present in the map, numbered in grid order; absent cells and the outside
 * of the grid carry the imposed value BND.  A direction of extent 1 carries no flux (a 2-D map (a,b) is the 3-D
 * layout (1,a,b)).  Parameters as diff3d.  The map travels in ipar:
 *   ip[3]=1 ip[4]=d1 ip[5]=d2 ip[6]=d3 ip[7]=unused ip[8]=n, then inv[n] (full-grid number, 1-based, of unknown c)
 *   and pres[d1*d2*d3] (new number of a full-grid cell, <= 0 if absent). */
#define RSM_DMAP_NPAR 8
#define RSM_DMAP_INV(ip) ((ip) + 9)
#define RSM_DMAP_PRES(ip) ((ip) + 9 + (ip)[8])
RSM_HD double rsm_dmap_value(const double *y, const int *pres, int inside, int cell, double bnd)
{
    if (!inside) return bnd;
    const int p = pres[cell];
    return (p > 0) ? y[p - 1] : bnd;
}

RSM_HD double rsm_diffmap_point(const double *y, int c, const int *inv, const int *pres, int d1, int d2, int d3,
                                const double *p)
{
    const double Dx = p[0], Dy = p[1], Dz = p[2], dx = p[3], dy = p[4], dz = p[5], r = p[6], bnd = p[7];
    const int s2 = d3, s1 = d2 * d3;
    const int cell = inv[c] - 1;
    const int j = cell / s1, rem = cell - j * s1, k = rem / s2, l = rem - k * s2;
    const double yc = y[c];
    double dY = -r * yc;
    if (d3 > 1)
        dY = dY - rsm_flux_div(rsm_dmap_value(y, pres, l > 0, cell - 1, bnd), yc,
                               rsm_dmap_value(y, pres, l < d3 - 1, cell + 1, bnd), Dx, dx);
    if (d2 > 1)
        dY = dY - rsm_flux_div(rsm_dmap_value(y, pres, k > 0, cell - s2, bnd), yc,
                               rsm_dmap_value(y, pres, k < d2 - 1, cell + s2, bnd), Dy, dy);
    if (d1 > 1)
        dY = dY - rsm_flux_div(rsm_dmap_value(y, pres, j > 0, cell - s1, bnd), yc,
                               rsm_dmap_value(y, pres, j < d1 - 1, cell + s1, bnd), Dz, dz);
    return dY;
}


/* Manufactured-solution variant of diff3d (SURVEY 8(c) protocol step 3): the diffusion-consumption model of
 * man/steady.3D.Rd plus a source  S = c0 + c1 * sj[j]*sk[k]*sl[l]  built from three 1-D tables of discrete sine
 * modes, sq[i] = sin(pi (i+1)/(dq+1)).  The product of the tables is an eigenvector of the 7-point operator with
 * imposed boundary value BND, so with c0 = r*BND and c1 = a*(lambda + r) the discrete steady state is
 * y* = BND + a*sj*sk*sl in closed form, at every grid size.  The tables arrive in rpar (after the nout outputs in
 * yout, src/steady_utils.c:109-124); p[0..7] as diff3d, p[8] = c0, p[9] = c1. */
#define RSM_D3MS_NPAR 10
RSM_HD double rsm_diff3dms_point(RSM_STATE y, int j, int k, int l, int d1, int d2, int d3, const rsm_d3_coef c,
                                 double c0, double c1, const double *tab)
{
    const double S = c0 + c1 * ((tab[j] * tab[d1 + k]) * tab[d1 + d2 + l]);
    return rsm_diff3d_point(y, j, k, l, d1, d2, d3, c, 0) + S;
}

/* The 2-D example of the vignette (vignettes/rootSolve.Rnw:615-663, `diffusion2D`): diffusion with ZERO-FLUX
 * boundaries, second-order consumption -r*C^2 and a constant production p in a handful of cells.  The only example
 * of the stodes path with a published run time (0.5 s / 0.16 s).  Without an imposed boundary value the Jacobian is
 * a bare Laplacian plus -2 r C on the diagonal.
 * p[0]=Dx p[1]=Dy p[2]=dx p[3]=dy p[4]=r p[5]=p ; ip[3]=1 ip[4]=d1 ip[5]=d2 ip[6]=1 ip[7]=nsrc, then nsrc 0-based
 * state-vector positions of the producing cells (the R code draws them with runif; the tests seed them). */
#define RSM_D2Q_NPAR 6
RSM_HD double rsm_diff2dq_point(RSM_STATE y, int j, int k, int d1, int d2, const double *p, const int *src, int nsrc)
{
    const double Dx = p[0], Dy = p[1], dx = p[2], dy = p[3], r = p[4], prod = p[5];
    const int idx = j * d2 + k;
    const double yc = y[idx];
    double dY = -r * yc * yc;
    for (int q = 0; q < nsrc; q++)
        if (src[q] == idx) { dY = dY + prod; break; }
    double flo = (k > 0) ? -Dx * (yc - y[idx - 1]) / dx : 0.0;
    double fhi = (k < d2 - 1) ? -Dx * (y[idx + 1] - yc) / dx : 0.0;
    dY = dY - (fhi - flo) / dx;
    flo = (j > 0) ? -Dy * (yc - y[idx - d2]) / dy : 0.0;
    fhi = (j < d1 - 1) ? -Dy * (y[idx + d2] - yc) / dy : 0.0;
    dY = dY - (fhi - flo) / dy;
    return dY;
}

/* The first example of man/stodes.Rd (:225-252): organic carbon in N sediment layers -- advection v, bioturbation
 * Kz (Kz/dx^2 = 1e4 against a decay of 5: diffusion dominated), imposed flux on top, zero gradient below, and a
 * NON-LOCAL exchange: a fraction of the top five layers is moved to layer N/2, which puts five entries far from the
 * diagonal into row N/2.  Solved by the reference with the default sparsetype = "sparseint".
 * p[0]=v p[1]=flux p[2]=dx p[3]=Kz p[4]=decay p[5]=ing ; ip[3]=1 ip[4]=N. */
#define RSM_SED_NPAR 6
RSM_HD double rsm_sediment_point(RSM_STATE y, int i, int N, const double *p)
{
    const double v = p[0], flux = p[1], dx = p[2], Kz = p[3], decay = p[4], ing = p[5];
    /* Flux through the upper (i) and lower (i+1) interface of layer i */
    const double up = (i > 0) ? v * y[i - 1] + -Kz * (y[i] - y[i - 1]) / dx : flux;
    const double dn = (i < N - 1) ? v * y[i] + -Kz * (y[i + 1] - y[i]) / dx : v * y[i] + -Kz * (y[i] - y[i]) / dx;
    double d = -(dn - up) / dx - decay * y[i];
    if (i < 5) d = d - ing * y[i];
    if (i == N / 2 - 1) d = d + ing * ((((y[0] + y[1]) + y[2]) + y[3]) + y[4]);
    return d;
}

#endif /* RSM_MODEL_MATH_H */
