// mg.cu -- aggregation multigrid on the grid the Jacobian comes from: the preconditioner of the Newton linear
// solve for steady.2D / steady.3D.  It takes the role the exact LDU factors play in the reference (nsfc/nnfc/nntc,
// src/sparse.f:1937-2563): make the step count of the solve independent of the conditioning of J.  Jacobi-
// preconditioned Krylov needs O(sqrt(cond)) steps -- 170 per solve on the 256^3 diffusion-reaction grid, and it
// stalls on a bare Laplacian; one V-cycle per step brings that to ~10, at any grid size.
//
//   levels     2x2x2 (2x2 in 2-D) piecewise-constant aggregation of CELLS, every species on its own: a coarse
//              unknown is (species, aggregate).  A dimension of extent <= 2 is not coarsened.  Coarsening stops
//              at <= 32 cells; that level is solved with an explicit dense inverse.
//   operators  Galerkin P'AP, which for this stencil is again a (2 nd + nspec)-point stencil on the coarse grid,
//              built by summing values (mg_build_coarse_kernel).  Entries that cross an aggregate face in a
//              coarsened direction are halved and the removed half is added to the diagonal: piecewise-constant
//              transfer makes the second-order part of P'AP twice too stiff (4 faces of size h^-2 against the
//              8 (2h)^-2 a coarse discretisation has), while the zeroth-order part (reaction, boundary terms =
//              the row sums) is summed exactly; the correction keeps the row sums and restores h-independent
//              convergence (scratch/mg_proto.py: 27 -> 9 CG steps at 64^3, zero-flux boundaries).
//   storage    every level, the finest included, holds its matrix by stencil slot ("ELL by diagonal"):
//              val[slot][row], slot 0 = diagonal, 1..6 = -j +j -k +k -l +l neighbour of the same species,
//              7.. = the other species of the same cell.  Columns are implicit, loads are coalesced, there is
//              no row pointer.  The finest level is a copy of the solver's row-wise arrays (one pass per Newton
//              iteration).  Matrices, Dinv and (one species) the d of the post-smoothing are single precision.
//   sharding   the top levels of a sharded solve are global: every rank works on its planes of one operator and
//              the halo planes of x / d travel before each application (NVLink stores + flags, mg_halo); below
//              them every rank continues on its own box.  RS_MG_GLOBAL_LEVELS (default 3), RS_MG_LOCAL.
//   smoother   Chebyshev polynomial of degree 2 in (block-)Jacobi-preconditioned form on [lmax/4, lmax], lmax
//              from a Gershgorin bound computed on the device; the cell block (nspec x nspec) is inverted
//              exactly, which takes the stiff reaction coupling out of the iteration.  Same smoother before and
//              after the coarse correction: for a symmetric matrix the cycle is a symmetric positive definite
//              operator and preconditions CG; BiCGStab uses the same cycle on nonsymmetric Jacobians.
//
// Bytes (level 0, N cells, one species, 7-point): a cycle makes 4 passes -- pre-smoothing 48 N, residual + restriction
// 45 N, the two post-smoothing passes 53 N and 61 N (matrix 28 N and Dinv 4 N in each, the rest vectors) -- 207 N in all
// (round-2 start: 4 x 56 N of matrix + ~170 N of vectors); levels below add 1/7 in 3-D.  The general form (several
// species, global levels of a sharded solve) takes 5 passes and stores r and d in double.  All kernels are thread-per-
// cell with grid-stride loops and run on the solver's stream without host synchronisation, so whole Krylov steps
// including the cycle replay from CUDA graphs.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <unordered_map>

#include "solver.cuh"

namespace rs {

constexpr int MG_MAXLEV = 28;
constexpr int MG_MAXS = 8;             // species handled (same bound as the block-Jacobi preconditioner)
constexpr int MG_COARSEST_CELLS = 32;  // stop coarsening here: <= 256 unknowns, inverted densely
constexpr int MG_MIN_CELLS = 64;       // smaller grids keep the (block-)Jacobi preconditioner
constexpr double MG_CHEB_RATIO = 4.0;  // smoothing interval [lmax / ratio, lmax]

// The level matrices are a preconditioner's data: stored in single precision (half the bytes of the four matrix passes
// of a cycle), promoted to double in every product.  The cycle stays a fixed linear operator; the Krylov method around
// it works on the double-precision Jacobian.  -DRS_MG_FP64 keeps them in double (A/B measurements).
#ifdef RS_MG_FP64
typedef double mg_real;
#else
typedef float mg_real;
#endif

// a vector addressed by (species, cell): internal vectors are packed (stride = cells of the level); the Krylov
// vectors at level 0 keep the solver's global indexing (stride = cells of the whole grid, off = first cell of the slab)
struct VecView {
    double *p;
    int stride, off;
    __device__ __forceinline__ double &at(int s, int c) const { return p[(long long)s * stride + off + c]; }
};

struct MgLevelDev {
    int nspec, d1, d2, d3, cyc1, cyc2, cyc3, ncell, n;
    int cb, ce;           // cells this rank works on: all of a local level's box, the rank's planes of a global level
    const mg_real *ell;   // [7 + nspec - 1][n]
    const mg_real *dinv;  // nspec == 1: 1/diag per row; else the inverse cell blocks [s][s'][cell]
    const double *lmax;   // Gershgorin bound of the (block-)Jacobi-preconditioned operator of this level
};

struct MgLevel {
    MgLevelDev dev;
    int f[3] = {1, 1, 1};   // coarsening factors towards the next level
    bool global = false;    // sharded solve: the level spans all ranks (global cell indices, halo planes exchanged)
    int cjoff = 0;          // transfer to the next level: plane 0 of that level's box in this level's aggregate numbering
    HaloPlan hp;            // global levels: the planes that travel before every application of the level's operator
    double *xp = nullptr, *dp = nullptr;   // x and d of the level (own buffers, or slices of the peer-mapped arena)
    double *x_lo = nullptr, *x_hi = nullptr, *d_lo = nullptr, *d_hi = nullptr;   // the slab neighbours' copies (halo push)
    DevBuf<mg_real> ell;
    DevBuf<mg_real> dinv;
    DevBuf<double> b, x, r, d;
    DevBuf<mg_real> df;     // one species, local level: the d of the post-smoothing (the only vector its two passes share)
};

struct Mg {
    int nlev = 0;
    MgLevel lev[MG_MAXLEV];
    DevBuf<double> lmax;      // one per level
    DevBuf<double> cinv;      // dense inverse of the coarsest operator, column-major
    DevBuf<double> cwork;     // 2 nc^2 work array of the inversion
    DevBuf<int> flag;         // != 0: the coarsest operator was singular to working precision
    int gstride = 0, goff = 0;   // level-0 view of the solver's vectors
    int nc = 0;               // unknowns of the coarsest level
    bool push = false;        // halo planes of the global levels are stored into the neighbours over NVLink
    DevBuf<unsigned> hticket;
    int nglobal = 0;          // number of global levels (0: single GPU, or a purely slab-local hierarchy)
    int tail = 0;             // first level of the one-launch tail (levels of <= MG_TAIL_CELLS cells); nlev: no tail
    DevBuf<unsigned char> tail_levels;   // MgTailLevel[nlev] on the device
};

// ---- stencil geometry ------------------------------------------------------------------------------------
struct Cell3 { int j, k, l; };

__device__ __forceinline__ Cell3 cell_coords(int c, int d2, int d3)
{
    Cell3 q;
    const int s1 = d2 * d3;
    q.j = c / s1;
    const int rem = c - q.j * s1;
    q.k = rem / d3;
    q.l = rem - q.k * d3;
    return q;
}

// cell index of the six neighbours (-j +j -k +k -l +l); a missing neighbour returns the cell itself (its matrix
// entry is zero) and has = false
__device__ __forceinline__ void neighbours(const MgLevelDev &L, int c, const Cell3 &q, int (&nb)[6], bool (&has)[6])
{
    const int s1 = L.d2 * L.d3;
    has[0] = q.j > 0 || (L.cyc1 && L.d1 > 1);          nb[0] = q.j > 0 ? c - s1 : c + (L.d1 - 1) * s1;
    has[1] = q.j < L.d1 - 1 || (L.cyc1 && L.d1 > 1);   nb[1] = q.j < L.d1 - 1 ? c + s1 : c - (L.d1 - 1) * s1;
    has[2] = q.k > 0 || (L.cyc2 && L.d2 > 1);          nb[2] = q.k > 0 ? c - L.d3 : c + (L.d2 - 1) * L.d3;
    has[3] = q.k < L.d2 - 1 || (L.cyc2 && L.d2 > 1);   nb[3] = q.k < L.d2 - 1 ? c + L.d3 : c - (L.d2 - 1) * L.d3;
    has[4] = q.l > 0 || (L.cyc3 && L.d3 > 1);          nb[4] = q.l > 0 ? c - 1 : c + (L.d3 - 1);
    has[5] = q.l < L.d3 - 1 || (L.cyc3 && L.d3 > 1);   nb[5] = q.l < L.d3 - 1 ? c + 1 : c - (L.d3 - 1);
#pragma unroll
    for (int t = 0; t < 6; t++)
        if (!has[t]) nb[t] = c;
}

#define MG_NSA (NS > 0 ? NS : MG_MAXS)

// (A v)(s, c) for every species s of cell c.  NS > 0: species count known at compile time.
template <int NS>
__device__ __forceinline__ void cell_apply(const MgLevelDev &L, int c, const int (&nb)[6], const VecView &v, double (&t)[MG_NSA])
{
    const int ns = NS > 0 ? NS : L.nspec;
    if constexpr (NS >= 1 && NS <= 3) {   // few species, known at compile time: every load of the cell before the first use
        constexpr int M = NS >= 1 ? NS : 1;
        const mg_real *__restrict__ ell = L.ell;
        mg_real a[M][6 + M];
        double vn[M][6], vcc[M];
#pragma unroll
        for (int s = 0; s < M; s++) {
            const long long r = (long long)s * L.ncell + c;
#pragma unroll
            for (int q = 0; q < 6 + M; q++) a[s][q] = ell[(long long)q * L.n + r];
#pragma unroll
            for (int q = 0; q < 6; q++) vn[s][q] = v.at(s, nb[q]);
            vcc[s] = v.at(s, c);
        }
#pragma unroll
        for (int s = 0; s < M; s++) {
            double acc = a[s][0] * vcc[s];
#pragma unroll
            for (int q = 0; q < 6; q++) acc += a[s][1 + q] * vn[s][q];
            int o = 0;
#pragma unroll
            for (int s2 = 0; s2 < M; s2++) {
                if (s2 == s) continue;
                acc += a[s][7 + o] * vcc[s2];
                o++;
            }
            t[s] = acc;
        }
        return;
    } else {
        double vc[MG_NSA];
#pragma unroll
        for (int s = 0; s < MG_NSA; s++)
            if (s < ns) vc[s] = v.at(s, c);
#pragma unroll
        for (int s = 0; s < MG_NSA; s++) {
            if (s >= ns) break;
            const long long r = (long long)s * L.ncell + c;
            double a = L.ell[r] * vc[s];
#pragma unroll
            for (int q = 0; q < 6; q++) a += L.ell[(long long)(1 + q) * L.n + r] * v.at(s, nb[q]);
            int o = 0;
#pragma unroll
            for (int s2 = 0; s2 < MG_NSA; s2++) {
                if (s2 >= ns) break;
                if (s2 == s) continue;
                a += L.ell[(long long)(7 + o) * L.n + r] * vc[s2];
                o++;
            }
            t[s] = a;
        }
    }
}

// w = Dinv v over the species of cell c (point inverse for one species, inverse cell block otherwise)
template <int NS>
__device__ __forceinline__ void cell_dinv(const MgLevelDev &L, int c, const double (&v)[MG_NSA], double (&w)[MG_NSA])
{
    const int ns = NS > 0 ? NS : L.nspec;
    if constexpr (NS == 1) {
        w[0] = L.dinv[c] * v[0];
    } else {
#pragma unroll
        for (int s = 0; s < MG_NSA; s++) {
            if (s >= ns) break;
            double a = 0.0;
#pragma unroll
            for (int s2 = 0; s2 < MG_NSA; s2++) {
                if (s2 >= ns) break;
                a += L.dinv[((long long)s * ns + s2) * L.ncell + c] * v[s2];
            }
            w[s] = a;
        }
    }
}

struct Cheb { double theta, inv_theta, c1, c2; };
__device__ __forceinline__ Cheb cheb_coeffs(const MgLevelDev &L)
{
    const double lmax = *L.lmax, lmin = lmax / MG_CHEB_RATIO;
    const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin);
    const double sigma = theta / delta, rho0 = 1.0 / sigma, rho1 = 1.0 / (2.0 * sigma - rho0);
    Cheb ch;
    ch.theta = theta;
    ch.inv_theta = 1.0 / theta;
    ch.c1 = rho1 * rho0;
    ch.c2 = 2.0 * rho1 / delta;
    return ch;
}

// every cycle kernel is a body (a __device__ function working on cells c0, c0 + cs, ...) with two callers: a __global__
// wrapper over the whole grid for the large levels, and the one-block tail kernel that runs all small levels in one launch
#define MG_CELL_LOOP(c, L) for (int c = (L).cb + c0; c < (L).ce; c += cs)
#define MG_GRID_ARGS (int)(blockIdx.x * blockDim.x + threadIdx.x), (int)(gridDim.x * blockDim.x)

// The same loop with the cell's coordinates carried along: two integer divisions before the loop (first cell and stride)
// instead of two per cell.
struct CellWalk {
    int c; Cell3 q; int cs; Cell3 sq; int d2, d3;
    __device__ __forceinline__ CellWalk(const MgLevelDev &L, int c0, int cs_) : c(L.cb + c0), cs(cs_), d2(L.d2), d3(L.d3)
    {
        q = cell_coords(c, d2, d3);
        sq = cell_coords(cs, d2, d3);
    }
    __device__ __forceinline__ void next()
    {
        c += cs;
        q.l += sq.l; if (q.l >= d3) { q.l -= d3; q.k++; }
        q.k += sq.k; if (q.k >= d2) { q.k -= d2; q.j++; }
        q.j += sq.j;
    }
};
#define MG_CELL_WALK(w, L) for (CellWalk w(L, c0, cs); w.c < (L).ce; w.next())

// ---- cycle kernels (thread per cell, all species of the cell) ------------------------------------------------
// first Chebyshev step from a zero guess:  d = Dinv b / theta   (x = d is implied)
template <int NS>
__device__ __forceinline__ void mg_first_body(const MgLevelDev &L, VecView b, VecView d, int c0, int cs)
{
    const int ns = NS > 0 ? NS : L.nspec;
    const Cheb ch = cheb_coeffs(L);
    MG_CELL_LOOP(c, L) {
        double bv[MG_NSA], w[MG_NSA];
#pragma unroll
        for (int s = 0; s < MG_NSA; s++)
            if (s < ns) bv[s] = b.at(s, c);
        cell_dinv<NS>(L, c, bv, w);
#pragma unroll
        for (int s = 0; s < MG_NSA; s++)
            if (s < ns) d.at(s, c) = ch.inv_theta * w[s];
    }
}

template <int NS>
__global__ void __launch_bounds__(256) mg_first_kernel(MgLevelDev L, VecView b, VecView d)
{
    mg_first_body<NS>(L, b, d, MG_GRID_ARGS);
}

// second Chebyshev step:  r1 = r - A d ;  d1 = c1 d + c2 Dinv r1 ;  xout = [xin +] d + d1
// DOT: also out[0] = <w, xout> (w = the right-hand side of the cycle: the <r, z> of preconditioned CG when this is
// the last kernel of the cycle)
template <int NS, bool DOT>
__device__ __forceinline__ double mg_second_body(const MgLevelDev &L, VecView r, VecView d, VecView xin, VecView xout, VecView w_,
                                                 int c0, int cs)
{
    const int ns = NS > 0 ? NS : L.nspec;
    const Cheb ch = cheb_coeffs(L);
    double acc[1] = {0.0};
    if constexpr (NS == 1) {   // one species: coordinates carried along, every load of the cell issued before the first use
        const double *dp = d.p + d.off;
        const double *rp = r.p + r.off;
        const mg_real *__restrict__ dinv = L.dinv;
        const mg_real *__restrict__ ell = L.ell;
        const double *xi = xin.p ? xin.p + xin.off : nullptr;
        MG_CELL_WALK(w, L) {
            const int c = w.c;
            int nb[6];
            bool has[6];
            neighbours(L, c, w.q, nb, has);
            mg_real a[7];
            double dn[6];
#pragma unroll
            for (int k = 0; k < 7; k++) a[k] = ell[(long long)k * L.n + c];
#pragma unroll
            for (int k = 0; k < 6; k++) dn[k] = dp[nb[k]];
            const double dv = dp[c], rc = rp[c], dic = dinv[c];
            const double xv = xi ? xi[c] : 0.0;
            const double wv = DOT ? w_.at(0, c) : 0.0;
            double t = a[0] * dv;
#pragma unroll
            for (int k = 0; k < 6; k++) t += a[1 + k] * dn[k];
            double xn = dv + (ch.c1 * dv + ch.c2 * (dic * (rc - t)));
            if (xi) xn += xv;
            xout.at(0, c) = xn;
            if (DOT) acc[0] += wv * xn;
        }
        return acc[0];
    } else {
        MG_CELL_WALK(cw, L) {
            const int c = cw.c;
            int nb[6];
            bool has[6];
            neighbours(L, c, cw.q, nb, has);
            double t[MG_NSA], r1[MG_NSA], w[MG_NSA];
            cell_apply<NS>(L, c, nb, d, t);
#pragma unroll
            for (int s = 0; s < MG_NSA; s++)
                if (s < ns) r1[s] = r.at(s, c) - t[s];
            cell_dinv<NS>(L, c, r1, w);
#pragma unroll
            for (int s = 0; s < MG_NSA; s++) {
                if (s >= ns) break;
                const double dv = d.at(s, c);
                double xn = dv + (ch.c1 * dv + ch.c2 * w[s]);
                if (xin.p != nullptr) xn += xin.at(s, c);
                xout.at(s, c) = xn;
                if (DOT) acc[0] += w_.at(s, c) * xn;
            }
        }
        return acc[0];
    }
}

template <int NS, bool DOT>
__global__ void __launch_bounds__(256, NS == 1 ? (DOT ? 3 : 4) : 1) mg_second_kernel(MgLevelDev L, VecView r, VecView d, VecView xin, VecView xout, VecView w_,
                                                        double *partials, unsigned *ticket, double *out, PeerRed pr)
{
    double acc[1];
    acc[0] = mg_second_body<NS, DOT>(L, r, d, xin, xout, w_, MG_GRID_ARGS);
    if (DOT) grid_reduce<1>(acc, 0u, partials, ticket, out, false, &pr);
}

// Both Chebyshev steps of the pre-smoothing in one pass (one species, local levels): the first step is pointwise,
// d = Dinv b / theta, so the neighbours' d are formed from their b and Dinv on the fly and d is never stored:
//     x = d + c1 d + c2 Dinv (b - A d)
__device__ __forceinline__ void mg_pre_body(const MgLevelDev &L, VecView b, VecView x, int c0, int cs)
{
    const Cheb ch = cheb_coeffs(L);
    const double *bp = b.p + b.off;
    const mg_real *__restrict__ dinv = L.dinv;
    const mg_real *__restrict__ ell = L.ell;
    MG_CELL_WALK(w, L) {
        const int c = w.c;
        int nb[6];
        bool has[6];
        neighbours(L, c, w.q, nb, has);
        // all 21 loads of the cell are issued before the first use (a missing neighbour reads the cell itself; its
        // matrix entry is zero): the pass is bound by memory latency, not by registers
        mg_real a[7];
        mg_real dn[6];
        double bn[6];
#pragma unroll
        for (int k = 0; k < 7; k++) a[k] = ell[(long long)k * L.n + c];
#pragma unroll
        for (int k = 0; k < 6; k++) { dn[k] = dinv[nb[k]]; bn[k] = bp[nb[k]]; }
        const double bc = bp[c], dic = dinv[c];
        const double dc = ch.inv_theta * dic * bc;
        double t = a[0] * dc;
#pragma unroll
        for (int k = 0; k < 6; k++) t += a[1 + k] * (ch.inv_theta * dn[k] * bn[k]);
        x.at(0, c) = dc + (ch.c1 * dc + ch.c2 * (dic * (bc - t)));
    }
}
template <int MINB>
__global__ void __launch_bounds__(256, MINB) mg_pre_kernel(MgLevelDev L, VecView b, VecView x)
{
    mg_pre_body(L, b, x, MG_GRID_ARGS);
}

// start of the post-smoothing:  r = b - A x ;  d = Dinv r / theta
template <int NS>
__device__ __forceinline__ void mg_resid_first_body(const MgLevelDev &L, VecView b, VecView x, VecView r, VecView d, int c0, int cs)
{
    const int ns = NS > 0 ? NS : L.nspec;
    const Cheb ch = cheb_coeffs(L);
    MG_CELL_WALK(cw, L) {
        const int c = cw.c;
        int nb[6];
        bool has[6];
        neighbours(L, c, cw.q, nb, has);
        double t[MG_NSA], rr[MG_NSA], w[MG_NSA];
        cell_apply<NS>(L, c, nb, x, t);
#pragma unroll
        for (int s = 0; s < MG_NSA; s++)
            if (s < ns) { rr[s] = b.at(s, c) - t[s]; r.at(s, c) = rr[s]; }
        cell_dinv<NS>(L, c, rr, w);
#pragma unroll
        for (int s = 0; s < MG_NSA; s++)
            if (s < ns) d.at(s, c) = ch.inv_theta * w[s];
    }
}

template <int NS>
__global__ void __launch_bounds__(256) mg_resid_first_kernel(MgLevelDev L, VecView b, VecView x, VecView r, VecView d)
{
    mg_resid_first_body<NS>(L, b, x, r, d, MG_GRID_ARGS);
}

// Post-smoothing with the prolongation folded in (one species, local levels), two passes that share one vector:
//     xt = x + P xc ;  r = b - A xt ;  d = Dinv r / theta                        (first pass: writes d)
//     xout = xt + d + c1 d + c2 Dinv (r - A d)                                   (second pass)
// The corrected value of a neighbour is formed on the fly from its x and its aggregate's xc, so xt is never stored
// (the second pass forms the cell's own again); r is not stored either: the second pass recovers it from d,
// r = theta a_cc d (Dinv = 1 / a_cc).  d is single precision like the matrix: preconditioner data.
struct AggMap {   // aggregate of a cell and of its six neighbours
    int s1, s2, s3, cd2, cd3;
    __device__ __forceinline__ AggMap(int f1, int f2, int f3, int cd2_, int cd3_)
        : s1(f1 == 2 ? 1 : 0), s2(f2 == 2 ? 1 : 0), s3(f3 == 2 ? 1 : 0), cd2(cd2_), cd3(cd3_) {}   // coarsening factors are 1 or 2: shifts
    __device__ __forceinline__ int own(const Cell3 &q) const { return ((q.j >> s1) * cd2 + (q.k >> s2)) * cd3 + (q.l >> s3); }
    // one coarse coordinate may differ (the neighbour's own coordinate with wrap-around); a missing neighbour maps to
    // an aggregate inside the grid
    __device__ __forceinline__ void around(const MgLevelDev &L, const Cell3 &q, int I, int (&In)[6]) const
    {
        const int J = q.j >> s1, K = q.k >> s2, Lc = q.l >> s3;
        const int jm = q.j > 0 ? q.j - 1 : L.d1 - 1, jp = q.j < L.d1 - 1 ? q.j + 1 : 0;
        const int km = q.k > 0 ? q.k - 1 : L.d2 - 1, kp = q.k < L.d2 - 1 ? q.k + 1 : 0;
        const int lm = q.l > 0 ? q.l - 1 : L.d3 - 1, lp = q.l < L.d3 - 1 ? q.l + 1 : 0;
        In[0] = I + ((jm >> s1) - J) * cd2 * cd3; In[1] = I + ((jp >> s1) - J) * cd2 * cd3;
        In[2] = I + ((km >> s2) - K) * cd3;       In[3] = I + ((kp >> s2) - K) * cd3;
        In[4] = I + ((lm >> s3) - Lc);            In[5] = I + ((lp >> s3) - Lc);
    }
};

__device__ __forceinline__ void mg_post_first_body(const MgLevelDev &L, int f1, int f2, int f3, int cd2, int cd3, VecView b, VecView x,
                                                   VecView xc, mg_real *df, int c0, int cs)
{
    const Cheb ch = cheb_coeffs(L);
    const AggMap agg(f1, f2, f3, cd2, cd3);
    const double *xp = x.p + x.off;
    const double *xcp = xc.p + xc.off;
    const mg_real *__restrict__ ell = L.ell;
    MG_CELL_WALK(w, L) {
        const int c = w.c;
        int nb[6], In[6];
        bool has[6];
        neighbours(L, c, w.q, nb, has);
        const int I = agg.own(w.q);
        agg.around(L, w.q, I, In);
        // every load before the first use; a missing neighbour reads the cell itself (zero matrix entry)
        mg_real a[7];
        double xn[6], cn[6];
#pragma unroll
        for (int k = 0; k < 7; k++) a[k] = ell[(long long)k * L.n + c];
#pragma unroll
        for (int k = 0; k < 6; k++) { xn[k] = xp[nb[k]]; cn[k] = xcp[In[k]]; }
        const double xv = xp[c], cv = xcp[I], bv = b.at(0, c), dic = L.dinv[c];
        double t = a[0] * (xv + cv);
#pragma unroll
        for (int k = 0; k < 6; k++) t += a[1 + k] * (xn[k] + cn[k]);
        df[c] = (mg_real)(ch.inv_theta * (dic * (bv - t)));
    }
}
template <int MINB>
__global__ void __launch_bounds__(256, MINB) mg_post_first_kernel(MgLevelDev L, int f1, int f2, int f3, int cd2, int cd3, VecView b, VecView x,
                                                               VecView xc, mg_real *df)
{
    mg_post_first_body(L, f1, f2, f3, cd2, cd3, b, x, xc, df, MG_GRID_ARGS);
}

// DOT: also <w, xout> (see mg_second_body).  xout may be x itself (every cell reads only its own x).
template <bool DOT>
__device__ __forceinline__ double mg_post_second_body(const MgLevelDev &L, int f1, int f2, int f3, int cd2, int cd3, VecView x, VecView xc,
                                                      const mg_real *df, VecView xout, VecView w_, int c0, int cs)
{
    const Cheb ch = cheb_coeffs(L);
    const AggMap agg(f1, f2, f3, cd2, cd3);
    const mg_real *__restrict__ ell = L.ell;
    double acc = 0.0;
    MG_CELL_WALK(w, L) {
        const int c = w.c;
        int nb[6];
        bool has[6];
        neighbours(L, c, w.q, nb, has);
        const int I = agg.own(w.q);
        mg_real a[7], dn[6];
#pragma unroll
        for (int k = 0; k < 7; k++) a[k] = ell[(long long)k * L.n + c];
#pragma unroll
        for (int k = 0; k < 6; k++) dn[k] = df[nb[k]];
        const double dv = df[c], xt = x.at(0, c) + xc.at(0, I);
        const double wv = DOT ? w_.at(0, c) : 0.0;
        const double a0 = a[0], dic = L.dinv[c];
        const double rc = ch.theta * dv * (a0 != 0.0 ? a0 : 1.0);   // dinv = 1 / a_cc (1 for an empty diagonal)
        double t = a0 * dv;
#pragma unroll
        for (int k = 0; k < 6; k++) t += a[1 + k] * (double)dn[k];
        const double xn = xt + (dv + (ch.c1 * dv + ch.c2 * (dic * (rc - t))));
        xout.at(0, c) = xn;
        if (DOT) acc += wv * xn;
    }
    return acc;
}
template <bool DOT, int MINB>
__global__ void __launch_bounds__(256, MINB) mg_post_second_kernel(MgLevelDev L, int f1, int f2, int f3, int cd2, int cd3, VecView x,
                                                                          VecView xc, const mg_real *df, VecView xout, VecView w_,
                                                                          double *partials, unsigned *ticket, double *out, PeerRed pr)
{
    double acc[1];
    acc[0] = mg_post_second_body<DOT>(L, f1, f2, f3, cd2, cd3, x, xc, df, xout, w_, MG_GRID_ARGS);
    if (DOT) grid_reduce<1>(acc, 0u, partials, ticket, out, false, &pr);
}

// residual of the fine level summed over each aggregate:  bc(s, I) = sum_{c in I} (b - A x)(s, c)
// thread per COARSE cell (cd1 x cd2 x cd3 of them); f: the coarsening factors
template <int NS>
__device__ __forceinline__ void mg_resid_restrict_body(const MgLevelDev &L, int f1, int f2, int f3, const MgLevelDev &Cc, int cjoff,
                                                       VecView b, VecView x, VecView bc, int c0, int cs)
{
    const int ns = NS > 0 ? NS : L.nspec;
    if constexpr (NS == 1) {   // one species: coordinates carried along; the (up to) two cells of a line that share the aggregate
                     // are staged together, every load before the first use
        const double *xp = x.p + x.off;
        const double *bp = b.p + b.off;
        const mg_real *__restrict__ ell = L.ell;
        MG_CELL_WALK(w, Cc) {
            const int Qj = w.q.j + cjoff;
            const int l0 = w.q.l * f3, nl = min(f3, L.d3 - l0);
            double sum = 0.0;
            for (int dj = 0; dj < f1; dj++) {
                const int j = Qj * f1 + dj;
                if (j >= L.d1) break;
                for (int dk = 0; dk < f2; dk++) {
                    const int k = w.q.k * f2 + dk;
                    if (k >= L.d2) break;
                    mg_real a[2][7];
                    double xn[2][6], xv[2], bv[2];
#pragma unroll
                    for (int dl = 0; dl < 2; dl++) {
                        Cell3 q;
                        q.j = j; q.k = k; q.l = l0 + (dl < nl ? dl : 0);   // an absent second cell re-reads the first
                        const int c = (j * L.d2 + k) * L.d3 + q.l;
                        int nb[6];
                        bool has[6];
                        neighbours(L, c, q, nb, has);
#pragma unroll
                        for (int t = 0; t < 7; t++) a[dl][t] = ell[(long long)t * L.n + c];
#pragma unroll
                        for (int t = 0; t < 6; t++) xn[dl][t] = xp[nb[t]];
                        xv[dl] = xp[c];
                        bv[dl] = bp[c];
                    }
#pragma unroll
                    for (int dl = 0; dl < 2; dl++) {
                        double t = a[dl][0] * xv[dl];
#pragma unroll
                        for (int q = 0; q < 6; q++) t += a[dl][1 + q] * xn[dl][q];
                        if (dl < nl) sum += bv[dl] - t;
                    }
                }
            }
            bc.at(0, w.c) = sum;
        }
        return;
    } else {
        MG_CELL_WALK(cw, Cc) {
            const int I = cw.c;
            Cell3 Q = cw.q;
            Q.j += cjoff;   // plane of the aggregate in the fine level's coordinates / f1
            double sum[MG_NSA];
#pragma unroll
            for (int s = 0; s < MG_NSA; s++) sum[s] = 0.0;
            for (int dj = 0; dj < f1; dj++) {
                const int j = Q.j * f1 + dj;
                if (j >= L.d1) break;
                for (int dk = 0; dk < f2; dk++) {
                    const int k = Q.k * f2 + dk;
                    if (k >= L.d2) break;
                    for (int dl = 0; dl < f3; dl++) {
                        const int l = Q.l * f3 + dl;
                        if (l >= L.d3) break;
                        const int c = (j * L.d2 + k) * L.d3 + l;
                        Cell3 q;
                        q.j = j; q.k = k; q.l = l;
                        int nb[6];
                        bool has[6];
                        neighbours(L, c, q, nb, has);
                        double t[MG_NSA];
                        cell_apply<NS>(L, c, nb, x, t);
#pragma unroll
                        for (int s = 0; s < MG_NSA; s++)
                            if (s < ns) sum[s] += b.at(s, c) - t[s];
                    }
                }
            }
#pragma unroll
            for (int s = 0; s < MG_NSA; s++)
                if (s < ns) bc.at(s, I) = sum[s];
        }
    }
}

template <int NS, int MINB>
__global__ void __launch_bounds__(256, MINB) mg_resid_restrict_kernel(MgLevelDev L, int f1, int f2, int f3, MgLevelDev Cc, int cjoff,
                                                                VecView b, VecView x, VecView bc)
{
    mg_resid_restrict_body<NS>(L, f1, f2, f3, Cc, cjoff, b, x, bc, MG_GRID_ARGS);
}

// x(s, c) += xc(s, aggregate of c)
__device__ __forceinline__ void mg_prolong_body(const MgLevelDev &L, int f1, int f2, int f3, int cd2, int cd3, int cjoff,
                                                VecView xc, VecView x, int c0, int cs)
{
    const int s1 = f1 == 2 ? 1 : 0, s2 = f2 == 2 ? 1 : 0, s3 = f3 == 2 ? 1 : 0;   // coarsening factors are 1 or 2
    MG_CELL_WALK(cw, L) {
        const int c = cw.c;
        const Cell3 q = cw.q;
        const int I = (((q.j >> s1) - cjoff) * cd2 + (q.k >> s2)) * cd3 + (q.l >> s3);
        for (int s = 0; s < L.nspec; s++) x.at(s, c) += xc.at(s, I);
    }
}

__global__ void __launch_bounds__(256) mg_prolong_kernel(MgLevelDev L, int f1, int f2, int f3, int cd2, int cd3, int cjoff,
                                                         VecView xc, VecView x)
{
    mg_prolong_body(L, f1, f2, f3, cd2, cd3, cjoff, xc, x, MG_GRID_ARGS);
}

// coarsest level:  x = Cinv b  (dense, column-major so that thread i reads a coalesced column slice)
__global__ void __launch_bounds__(256) mg_coarse_solve_kernel(int nc, const double *__restrict__ cinv, const double *__restrict__ b,
                                                              double *__restrict__ x)
{
    extern __shared__ double sb[];
    for (int i = threadIdx.x; i < nc; i += blockDim.x) sb[i] = b[i];
    __syncthreads();
    for (int i = threadIdx.x; i < nc; i += blockDim.x) {
        double a = 0.0;
        for (int j = 0; j < nc; j++) a += cinv[(size_t)j * nc + i] * sb[j];
        x[i] = a;
    }
}

// ---- halo planes of a global level over NVLink ----------------------------------------------------------------
// The rank's first / last plane of every species goes straight into the slab neighbours' copy of the vector (peer
// stores), the last block raises the neighbours' flags and waits for theirs, so everything later in the stream sees
// complete halos.  One small launch instead of an NCCL send/recv group: a cycle has twelve of these.  (Pushing from the producing kernel and waiting in the consuming one, as the
// Krylov kernels do for their one vector, was measured SLOWER here -- 35.7 against 30.7 ms at 2 GPUs: every block of
// the large level kernels then pays the system-scope fence and the flag wait that these two small kernels pay once.)
__global__ void __launch_bounds__(256) mg_halo_push_kernel(int nspec, int ncell, int plane, int j0, int j1, const double *__restrict__ v,
                                                           HaloPush hp, unsigned *ticket)
{
    const long long per = (long long)nspec * plane;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < 2 * per; q += (long long)gridDim.x * blockDim.x) {
        const int side = q >= per ? 1 : 0;
        const long long w = q - side * per;
        const int s = (int)(w / plane), o = (int)(w - (long long)s * plane);
        const long long i = (long long)s * ncell + (long long)(side ? j1 - 1 : j0) * plane + o;
        double *dst = side ? hp.hi : hp.lo;
        if (dst != nullptr) dst[i] = v[i];
    }
    // the last block to finish raises the neighbours' flags -- and then stays until the neighbours' own flags for this
    // exchange have arrived: the kernel ends when the halos are complete, so one launch does what the push kernel and
    // a separate waiting kernel did (raising never depends on waiting: no rank can block another)
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(ticket, 1u);
        if (t == gridDim.x - 1) {
            *ticket = 0;
            const unsigned seq = *hp.sig.seq + 1u;
            __threadfence_system();
            if (hp.sig.tell_lo) *(volatile double *)hp.sig.tell_lo = (double)seq;
            if (hp.sig.tell_hi) *(volatile double *)hp.sig.tell_hi = (double)seq;
            *hp.sig.seq = seq;
            const double want = (double)seq;
            const long long t0 = clock64();
            if (hp.sig.wait_lo) while (*(volatile const double *)hp.sig.wait_lo < want) { if (clock64() - t0 > 4000000000LL) __trap(); }
            if (hp.sig.wait_hi) while (*(volatile const double *)hp.sig.wait_hi < want) { if (clock64() - t0 > 4000000000LL) __trap(); }
            __threadfence_system();
        }
    }
}

// A flag of the exchange above says "my planes have arrived", not "I have finished reading yours".  Two consecutive
// exchanges target different buffers (d, x, then the next level's) except the x of the deepest global level: pushed
// before its restriction and again after its prolongation, with only this rank's local levels -- no exchange -- in
// between.  There the reader acknowledges: after the restriction it raises a flag of its own ("done reading"), and the
// second push waits for the neighbours' before it writes into their halo planes.  Two one-thread kernels per cycle;
// the wait is over long before it is asked for (the neighbour's whole local hierarchy lies in between).
__global__ void mg_halo_raise_kernel(HaloSignal sig)
{
    const unsigned seq = *sig.seq + 1u;
    __threadfence_system();
    if (sig.tell_lo) *(volatile double *)sig.tell_lo = (double)seq;
    if (sig.tell_hi) *(volatile double *)sig.tell_hi = (double)seq;
    *sig.seq = seq;
}
__global__ void mg_halo_wait_kernel(HaloSignal sig) { halo_wait(sig); }

// ---- the small levels in one launch ---------------------------------------------------------------------------
// Below a few thousand cells a level is a handful of microseconds of work per kernel and six kernels per visit: launch
// latency, not bandwidth.  One block walks all those levels -- down, the dense coarsest solve, up -- with a block
// barrier where the separate kernels had a launch boundary.  When the whole grid is that small (config C1) the
// entire cycle is this one launch, its <r, z> included.
struct MgTailLevel {
    MgLevelDev D;
    int f1, f2, f3;
    double *b, *x, *r, *d;
    mg_real *df;
};
constexpr int MG_TAIL_THREADS = 1024;
constexpr int MG_TAIL_CELLS = 4096;

template <int NS>
__global__ void __launch_bounds__(MG_TAIL_THREADS) mg_tail_kernel(const MgTailLevel *__restrict__ lv, int l0, int nlev, int nc,
                                                                  const double *__restrict__ cinv, VecView b0, VecView xout0,
                                                                  int want_dot, double *dot_out)
{
    const int c0 = threadIdx.x, cs = blockDim.x;
    const VecView none = {nullptr, 0, 0};
    for (int l = l0; l < nlev - 1; l++) {   // down
        const MgTailLevel T = lv[l];
        const MgLevelDev Dn = lv[l + 1].D;
        const VecView b = (l == l0 && b0.p != nullptr) ? b0 : VecView{T.b, T.D.ncell, 0};
        const VecView x = {T.x, T.D.ncell, 0}, d = {T.d, T.D.ncell, 0};
        if (NS == 1) {
            mg_pre_body(T.D, b, x, c0, cs);
        } else {
            mg_first_body<NS>(T.D, b, d, c0, cs);
            __syncthreads();
            mg_second_body<NS, false>(T.D, b, d, none, x, none, c0, cs);
        }
        __syncthreads();
        mg_resid_restrict_body<NS>(T.D, T.f1, T.f2, T.f3, Dn, 0, b, x, VecView{lv[l + 1].b, Dn.ncell, 0}, c0, cs);
        __syncthreads();
    }
    {   // coarsest level: x = Cinv b
        const MgTailLevel T = lv[nlev - 1];
        const double *bb = T.b;
        for (int i = c0; i < nc; i += cs) {
            double a = 0.0;
            for (int j = 0; j < nc; j++) a += cinv[(size_t)j * nc + i] * bb[j];
            T.x[i] = a;
        }
        __syncthreads();
    }
    double acc = 0.0;
    for (int l = nlev - 2; l >= l0; l--) {   // up
        const MgTailLevel T = lv[l];
        const MgLevelDev Dn = lv[l + 1].D;
        const bool top = l == l0 && b0.p != nullptr;
        const VecView b = top ? b0 : VecView{T.b, T.D.ncell, 0};
        const VecView x = {T.x, T.D.ncell, 0}, r = {T.r, T.D.ncell, 0}, d = {T.d, T.D.ncell, 0};
        const VecView xc = {lv[l + 1].x, Dn.ncell, 0};
        if (NS == 1) {
            mg_post_first_body(T.D, T.f1, T.f2, T.f3, Dn.d2, Dn.d3, b, x, xc, T.df, c0, cs);
            __syncthreads();
            if (top && want_dot) acc = mg_post_second_body<true>(T.D, T.f1, T.f2, T.f3, Dn.d2, Dn.d3, x, xc, T.df, xout0, b, c0, cs);
            else mg_post_second_body<false>(T.D, T.f1, T.f2, T.f3, Dn.d2, Dn.d3, x, xc, T.df, top ? xout0 : x, none, c0, cs);
        } else {
            mg_prolong_body(T.D, T.f1, T.f2, T.f3, Dn.d2, Dn.d3, 0, xc, x, c0, cs);
            __syncthreads();
            mg_resid_first_body<NS>(T.D, b, x, r, d, c0, cs);
            __syncthreads();
            if (top && want_dot) acc = mg_second_body<NS, true>(T.D, r, d, x, xout0, b, c0, cs);
            else mg_second_body<NS, false>(T.D, r, d, x, top ? xout0 : x, none, c0, cs);
        }
        __syncthreads();
    }
    if (want_dot) {   // <r, z> over the block
        __shared__ double red[32];
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x < 32) {
            double v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
            v = warp_sum(v);
            if (threadIdx.x == 0) dot_out[0] = v;
        }
    }
}

// ---- set-up kernels ------------------------------------------------------------------------------------------
__device__ __forceinline__ void atomic_max_pos(double *addr, double v)   // v >= 0: the bit patterns order like the values
{
    atomicMax(reinterpret_cast<unsigned long long *>(addr), (unsigned long long)__double_as_longlong(v));
}

// Level 0 from the solver's row-wise arrays (global indices, columns ascending).  Thread per (species, local cell).
// An entry of the same species goes to the slot of the neighbour it addresses, an entry of another species of the
// same cell to that species' slot; anything else -- a column outside the rank's slab, a pattern entry that is not
// part of the stencil -- is left out of the preconditioner.  ell must be zero on entry.
__global__ void __launch_bounds__(256) mg_build_l0_kernel(MgLevelDev L, int gstride, int goff, const int *__restrict__ rowptr,
                                                          const int *__restrict__ col, const double *__restrict__ val,
                                                          mg_real *__restrict__ ell)
{
    const int own = L.ce - L.cb;
    const long long total = (long long)L.nspec * own;
    for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (long long)gridDim.x * blockDim.x) {
        const int s = (int)(w / own), c = L.cb + (int)(w - (long long)s * own);
        const long long r = (long long)s * L.ncell + c;
        const Cell3 q = cell_coords(c, L.d2, L.d3);
        int nb[6];
        bool has[6];
        neighbours(L, c, q, nb, has);
        const long long g = (long long)s * gstride + goff + c;
        double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        for (int p = rowptr[g]; p < rowptr[g + 1]; p++) {
            const int cj = col[p];
            const int s2 = cj / gstride;
            const int c2 = cj - s2 * gstride - goff;
            if (c2 < 0 || c2 >= L.ncell) continue;
            const double v = val[p];
            if (s2 != s) {
                if (c2 == c) ell[(long long)(7 + (s2 < s ? s2 : s2 - 1)) * L.n + r] += (mg_real)v;
                continue;
            }
            if (c2 == c) { acc[0] += v; continue; }
#pragma unroll
            for (int t = 0; t < 6; t++)
                if (has[t] && nb[t] == c2) { acc[1 + t] += v; break; }
        }
#pragma unroll
        for (int t = 0; t < 7; t++) ell[(long long)t * L.n + r] = (mg_real)acc[t];
    }
}

// One species: the smoother data of a row whose seven entries are in registers (what mg_smoother_setup_kernel reads back
// from the level for the general case): dinv = 1 / a_cc of the stored (rounded) value, and the row's Gershgorin bound.
__device__ __forceinline__ double row_smoother_data(const mg_real (&e)[7], mg_real *dinv_out)
{
    const double a = (double)e[0];
    const double di = fabs(a) > 0.0 ? 1.0 / a : 1.0;
    double ts = 0.0;
#pragma unroll
    for (int t = 0; t < 6; t++) ts += fabs((double)e[1 + t]);
    *dinv_out = (mg_real)di;
    return 1.0 + fabs(di) * ts;
}

// Level 0, one species, pattern verified as the grid's own (grid_stencil.cuh): the entries of a row are its neighbours
// in ascending column order -j -k -l diag +l +k +j, the present ones only -- no column is read, nothing is searched.
// Writes all seven slots (no zero-fill needed), dinv and the level's Gershgorin bound.
__global__ void __launch_bounds__(256, 4) mg_build_l0_grid_kernel(MgLevelDev L, const int *__restrict__ rowptr, const double *__restrict__ val,
                                                                  mg_real *__restrict__ ell, mg_real *__restrict__ dinv, double *__restrict__ lmax)
{
    const int c0 = blockIdx.x * blockDim.x + threadIdx.x, cs = gridDim.x * blockDim.x;
    double worst = 0.0;
    MG_CELL_WALK(w, L) {
        const int c = w.c;
        const Cell3 q = w.q;
        // slot of each candidate in column order, and whether the row has it
        const bool has[7] = {q.j > 0, q.k > 0, q.l > 0, true, q.l < L.d3 - 1, q.k < L.d2 - 1, q.j < L.d1 - 1};
        constexpr int slot[7] = {1, 3, 5, 0, 6, 4, 2};
        const int p0 = rowptr[c];
        double v[7];
        int p = p0;
#pragma unroll
        for (int t = 0; t < 7; t++) {
            v[t] = has[t] ? val[p] : 0.0;
            p += has[t] ? 1 : 0;
        }
        mg_real e[7];
#pragma unroll
        for (int t = 0; t < 7; t++) e[slot[t]] = (mg_real)v[t];
#pragma unroll
        for (int t = 0; t < 7; t++) ell[(long long)t * L.n + c] = e[t];
        mg_real di;
        worst = fmax(worst, row_smoother_data(e, &di));
        dinv[c] = di;
    }
    worst = warp_max(worst);
    if ((threadIdx.x & 31) == 0 && worst > 0.0) atomic_max_pos(lmax, worst);
}

// Galerkin coarse operator, one species: as mg_build_coarse_kernel below, with the coordinates carried along, the (up
// to) two cells of a line that share the aggregate loaded together, no division, and the coarse level's smoother data
// formed from the sums while they are in registers.
__global__ void __launch_bounds__(256, 4) mg_build_coarse1_kernel(MgLevelDev F, MgLevelDev Cc, int f1, int f2, int f3, int cjoff, double w,
                                                                  mg_real *__restrict__ ellc, mg_real *__restrict__ dinvc,
                                                                  double *__restrict__ lmaxc)
{
    const int c0 = blockIdx.x * blockDim.x + threadIdx.x, cs = gridDim.x * blockDim.x;
    const mg_real *__restrict__ ellf = F.ell;
    double worst = 0.0;
    MG_CELL_WALK(cw, Cc) {
        const Cell3 Q = cw.q;
        double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        const int l0 = Q.l * f3, nl = min(f3, F.d3 - l0);
        for (int dj = 0; dj < f1; dj++) {
            const int j = (Q.j + cjoff) * f1 + dj;
            if (j >= F.d1) break;
            for (int dk = 0; dk < f2; dk++) {
                const int k = Q.k * f2 + dk;
                if (k >= F.d2) break;
                mg_real a[2][7];
#pragma unroll
                for (int dl = 0; dl < 2; dl++) {
                    const int c = (j * F.d2 + k) * F.d3 + l0 + (dl < nl ? dl : 0);   // an absent second cell re-reads the first
#pragma unroll
                    for (int t = 0; t < 7; t++) a[dl][t] = ellf[(long long)t * F.n + c];
                }
#pragma unroll
                for (int dl = 0; dl < 2; dl++) {
                    if (dl >= nl) continue;
                    const int l = l0 + dl;
                    acc[0] += (double)a[dl][0];
                    // neighbour t of the fine cell: its aggregate's coordinate in the direction of t (wrap-around as in
                    // neighbours()); present iff the grid goes on there or is cyclic
                    const int co[3] = {j, k, l}, dsz[3] = {F.d1, F.d2, F.d3}, cyc[3] = {F.cyc1, F.cyc2, F.cyc3};
                    const int fq[3] = {f1, f2, f3}, IQ[3] = {Q.j, Q.k, Q.l};
#pragma unroll
                    for (int t = 0; t < 6; t++) {
                        const int dim = t >> 1;
                        const bool up = t & 1;
                        const bool inside = up ? co[dim] < dsz[dim] - 1 : co[dim] > 0;
                        if (!inside && !(cyc[dim] && dsz[dim] > 1)) continue;
                        const int cn = up ? (inside ? co[dim] + 1 : 0) : (inside ? co[dim] - 1 : dsz[dim] - 1);
                        const double v = (double)a[dl][1 + t];
                        const int an = (fq[dim] == 2 ? cn >> 1 : cn) - (dim == 0 ? cjoff : 0);
                        if (dim == 0 && (an < 0 || an >= Cc.d1)) continue;        // another rank's
                        if (an == IQ[dim]) acc[0] += v;                           // inside the aggregate
                        else if (fq[dim] == 2) { acc[1 + t] += w * v; acc[0] += (1.0 - w) * v; }
                        else acc[1 + t] += v;
                    }
                }
            }
        }
        mg_real e[7];
#pragma unroll
        for (int t = 0; t < 7; t++) { e[t] = (mg_real)acc[t]; ellc[(long long)t * Cc.n + cw.c] = e[t]; }
        mg_real di;
        worst = fmax(worst, row_smoother_data(e, &di));
        dinvc[cw.c] = di;
    }
    worst = warp_max(worst);
    if ((threadIdx.x & 31) == 0 && worst > 0.0) atomic_max_pos(lmaxc, worst);
}

// Galerkin coarse operator by summation.  Thread per (species, coarse cell).  w = weight kept by an entry that
// crosses an aggregate face in a coarsened direction (the rest goes to the diagonal, see the file header).
__global__ void __launch_bounds__(256) mg_build_coarse_kernel(MgLevelDev F, MgLevelDev Cc, int f1, int f2, int f3, int cjoff, double w,
                                                              mg_real *__restrict__ ellc)
{
    // cjoff: plane of the coarse box's plane 0 in the fine level's aggregate numbering (the fine level is global,
    // the coarse one the rank's own box); an aggregate outside the coarse box belongs to another rank: dropped
    const int own = Cc.ce - Cc.cb;
    const long long total = (long long)Cc.nspec * own;
    const int nother = F.nspec - 1;
    for (long long W = (long long)blockIdx.x * blockDim.x + threadIdx.x; W < total; W += (long long)gridDim.x * blockDim.x) {
        const int s = (int)(W / own), I = Cc.cb + (int)(W - (long long)s * own);
        const long long R = (long long)s * Cc.ncell + I;
        const Cell3 Q = cell_coords(I, Cc.d2, Cc.d3);
        double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        double oth[MG_MAXS];
        for (int o = 0; o < MG_MAXS; o++) oth[o] = 0.0;
        const int fq[3] = {f1, f2, f3};
        const int IQ[3] = {Q.j, Q.k, Q.l};
        for (int dj = 0; dj < f1; dj++) {
            const int j = (Q.j + cjoff) * f1 + dj;
            if (j >= F.d1) break;
            for (int dk = 0; dk < f2; dk++) {
                const int k = Q.k * f2 + dk;
                if (k >= F.d2) break;
                for (int dl = 0; dl < f3; dl++) {
                    const int l = Q.l * f3 + dl;
                    if (l >= F.d3) break;
                    const int c = (j * F.d2 + k) * F.d3 + l;
                    const long long r = (long long)s * F.ncell + c;
                    Cell3 q;
                    q.j = j; q.k = k; q.l = l;
                    int nb[6];
                    bool has[6];
                    neighbours(F, c, q, nb, has);
                    acc[0] += F.ell[r];
#pragma unroll
                    for (int t = 0; t < 6; t++) {
                        if (!has[t]) continue;
                        const double v = F.ell[(long long)(1 + t) * F.n + r];
                        const Cell3 qn = cell_coords(nb[t], F.d2, F.d3);
                        const int dim = t >> 1;
                        const int cn = dim == 0 ? qn.j : dim == 1 ? qn.k : qn.l;
                        const int an = cn / fq[dim] - (dim == 0 ? cjoff : 0);     // the neighbour's aggregate, coarse coordinates
                        if (dim == 0 && (an < 0 || an >= Cc.d1)) continue;        // another rank's
                        if (an == IQ[dim]) acc[0] += v;                           // inside the aggregate
                        else if (fq[dim] == 2) { acc[1 + t] += w * v; acc[0] += (1.0 - w) * v; }
                        else acc[1 + t] += v;
                    }
                    for (int o = 0; o < nother; o++) oth[o] += F.ell[(long long)(7 + o) * F.n + r];
                }
            }
        }
#pragma unroll
        for (int t = 0; t < 7; t++) ellc[(long long)t * Cc.n + R] = (mg_real)acc[t];
        for (int o = 0; o < nother; o++) ellc[(long long)(7 + o) * Cc.n + R] = (mg_real)oth[o];
    }
}

// Smoother data of a level: the inverse of every cell block (nspec = 1: of the diagonal) and the Gershgorin bound
//   lmax = 1 + max over rows of  sum_s' |Binv(s,s')| * (sum over the six neighbour entries of row (s', cell))
// of the block-Jacobi-preconditioned operator Binv A = I + Binv T.  lmax[0] must be zero on entry.
__global__ void __launch_bounds__(128) mg_smoother_setup_kernel(MgLevelDev L, mg_real *__restrict__ dinv, double *__restrict__ lmax)
{
    const int ns = L.nspec;
    double worst = 0.0;
    const int c0 = blockIdx.x * blockDim.x + threadIdx.x, cs = gridDim.x * blockDim.x;
    MG_CELL_LOOP(c, L) {
        double A[MG_MAXS][MG_MAXS], B[MG_MAXS][MG_MAXS], tsum[MG_MAXS];
        for (int s = 0; s < ns; s++) {
            const long long r = (long long)s * L.ncell + c;
            int o = 0;
            for (int s2 = 0; s2 < ns; s2++) {
                A[s][s2] = (s2 == s) ? L.ell[r] : L.ell[(long long)(7 + o++) * L.n + r];
                B[s][s2] = (s == s2) ? 1.0 : 0.0;
            }
            double ts = 0.0;
            for (int t = 0; t < 6; t++) ts += fabs(L.ell[(long long)(1 + t) * L.n + r]);
            tsum[s] = ts;
        }
        bool singular = false;
        for (int k = 0; k < ns && !singular; k++) {   // Gauss-Jordan with partial pivoting
            int piv = k;
            double best = fabs(A[k][k]);
            for (int s = k + 1; s < ns; s++)
                if (fabs(A[s][k]) > best) { best = fabs(A[s][k]); piv = s; }
            if (!(best > 0.0)) { singular = true; break; }
            if (piv != k)
                for (int q = 0; q < ns; q++) {
                    const double ta = A[k][q]; A[k][q] = A[piv][q]; A[piv][q] = ta;
                    const double tb = B[k][q]; B[k][q] = B[piv][q]; B[piv][q] = tb;
                }
            const double inv = 1.0 / A[k][k];
            for (int q = 0; q < ns; q++) { A[k][q] *= inv; B[k][q] *= inv; }
            for (int s = 0; s < ns; s++) {
                if (s == k) continue;
                const double f = A[s][k];
                if (f == 0.0) continue;
                for (int q = 0; q < ns; q++) { A[s][q] -= f * A[k][q]; B[s][q] -= f * B[k][q]; }
            }
        }
        if (singular)
            for (int s = 0; s < ns; s++)
                for (int q = 0; q < ns; q++) B[s][q] = (s == q) ? 1.0 : 0.0;
        for (int s = 0; s < ns; s++) {
            double g = 0.0;
            for (int q = 0; q < ns; q++) {
                g += fabs(B[s][q]) * tsum[q];
                if (ns == 1) dinv[c] = (mg_real)B[0][0];
                else dinv[((long long)s * ns + q) * L.ncell + c] = (mg_real)B[s][q];
            }
            worst = fmax(worst, 1.0 + g);
        }
    }
    worst = warp_max(worst);
    if ((threadIdx.x & 31) == 0 && worst > 0.0) atomic_max_pos(lmax, worst);
}

// out = { max |a_ij - a_ji| , max |a_ij| } over the rank's rows of level 0, read off the slot-major copy: the transposed
// entry of slot t of cell c is slot t^1 of the neighbour (the other species' slot of the same cell) -- no search.
// Pairs whose partner row belongs to another rank are skipped.  Single precision is enough for the 1e-6 threshold of
// the symmetric / nonsymmetric decision (krylov.cu).
__global__ void __launch_bounds__(256, 4) mg_asymmetry_kernel(MgLevelDev L, double *partials, unsigned *ticket, double *out)
{
    const int c0 = blockIdx.x * blockDim.x + threadIdx.x, cs = gridDim.x * blockDim.x;
    double acc[2] = {0.0, 0.0};
    if (L.nspec == 1) {   // coordinates carried along, the thirteen loads of a cell issued together
        const mg_real *__restrict__ ell = L.ell;
        MG_CELL_WALK(w, L) {
            const int c = w.c;
            int nb[6];
            bool has[6];
            neighbours(L, c, w.q, nb, has);
            mg_real a[7], tr[6];
#pragma unroll
            for (int t = 0; t < 7; t++) a[t] = ell[(long long)t * L.n + c];
#pragma unroll
            for (int t = 0; t < 6; t++) tr[t] = ell[(long long)(1 + (t ^ 1)) * L.n + nb[t]];   // a missing neighbour: the cell itself
            float hi = fabsf(a[0]), df = 0.f;
#pragma unroll
            for (int t = 0; t < 6; t++) {
                hi = fmaxf(hi, fabsf(a[1 + t]));
                if (has[t] && nb[t] >= L.cb && nb[t] < L.ce) df = fmaxf(df, fabsf(a[1 + t] - tr[t]));
            }
            acc[0] = fmax(acc[0], (double)df);
            acc[1] = fmax(acc[1], (double)hi);
        }
        grid_reduce<2>(acc, 3u, partials, ticket, out, false);
        return;
    }
    MG_CELL_LOOP(c, L) {
        const Cell3 q = cell_coords(c, L.d2, L.d3);
        int nb[6];
        bool has[6];
        neighbours(L, c, q, nb, has);
        for (int s = 0; s < L.nspec; s++) {
            const long long r = (long long)s * L.ncell + c;
            acc[1] = fmax(acc[1], fabs((double)L.ell[r]));
#pragma unroll
            for (int t = 0; t < 6; t++) {
                const double v = L.ell[(long long)(1 + t) * L.n + r];
                acc[1] = fmax(acc[1], fabs(v));
                if (has[t] && nb[t] >= L.cb && nb[t] < L.ce)
                    acc[0] = fmax(acc[0], fabs(v - (double)L.ell[(long long)(1 + (t ^ 1)) * L.n + (long long)s * L.ncell + nb[t]]));
            }
            int o = 0;
            for (int s2 = 0; s2 < L.nspec; s2++) {
                if (s2 == s) continue;
                const double v = L.ell[(long long)(7 + o) * L.n + r];
                const int o2 = s < s2 ? s : s - 1;   // slot of species s in the row of species s2
                acc[1] = fmax(acc[1], fabs(v));
                acc[0] = fmax(acc[0], fabs(v - (double)L.ell[(long long)(7 + o2) * L.n + (long long)s2 * L.ncell + c]));
                o++;
            }
        }
    }
    grid_reduce<2>(acc, 3u, partials, ticket, out, false);
}

// Dense inverse of the coarsest operator (n <= 256) by Gauss-Jordan with partial pivoting, one block.
// work: 2 n^2 doubles (A | Inv), row-major; cinv: the inverse, column-major.  flag: set when a pivot vanishes.
__global__ void __launch_bounds__(1024) mg_coarse_invert_kernel(MgLevelDev L, double *__restrict__ work, double *__restrict__ cinv,
                                                                int *__restrict__ flag)
{
    const int n = L.n, tid = threadIdx.x, nt = blockDim.x;
    double *A = work, *B = work + (size_t)n * n;
    __shared__ double fcol[256];
    __shared__ int s_piv;
    __shared__ double s_inv;
    for (int e = tid; e < n * n; e += nt) { A[e] = 0.0; B[e] = (e / n == e % n) ? 1.0 : 0.0; }
    __syncthreads();
    for (int r = tid; r < n; r += nt) {   // expand the stencil rows
        const int s = r / L.ncell, c = r - s * L.ncell;
        const Cell3 q = cell_coords(c, L.d2, L.d3);
        int nb[6];
        bool has[6];
        neighbours(L, c, q, nb, has);
        A[(size_t)r * n + r] += L.ell[r];
        for (int t = 0; t < 6; t++)
            if (has[t]) A[(size_t)r * n + s * L.ncell + nb[t]] += L.ell[(size_t)(1 + t) * n + r];
        int o = 0;
        for (int s2 = 0; s2 < L.nspec; s2++) {
            if (s2 == s) continue;
            A[(size_t)r * n + s2 * L.ncell + c] += L.ell[(size_t)(7 + o) * n + r];
            o++;
        }
    }
    __syncthreads();
    for (int k = 0; k < n; k++) {
        if (tid == 0) {
            int piv = k;
            double best = fabs(A[(size_t)k * n + k]);
            for (int i = k + 1; i < n; i++) {
                const double v = fabs(A[(size_t)i * n + k]);
                if (v > best) { best = v; piv = i; }
            }
            s_piv = piv;
            if (!(best > 1.0e-300)) { *flag = 1; s_inv = 1.0; }
            else s_inv = 1.0 / A[(size_t)piv * n + k];
        }
        __syncthreads();
        const int piv = s_piv;
        const double inv = s_inv;
        for (int j = tid; j < n; j += nt) {   // swap rows k <-> piv and scale the pivot row
            const double a = A[(size_t)piv * n + j], b = B[(size_t)piv * n + j];
            if (piv != k) {
                A[(size_t)piv * n + j] = A[(size_t)k * n + j];
                B[(size_t)piv * n + j] = B[(size_t)k * n + j];
            }
            A[(size_t)k * n + j] = a * inv;
            B[(size_t)k * n + j] = b * inv;
        }
        __syncthreads();
        for (int i = tid; i < n; i += nt) fcol[i] = (i == k) ? 0.0 : A[(size_t)i * n + k];
        __syncthreads();
        for (int e = tid; e < n * n; e += nt) {
            const int i = e / n, j = e - i * n;
            const double f = fcol[i];
            if (f != 0.0) {
                A[e] -= f * A[(size_t)k * n + j];
                B[e] -= f * B[(size_t)k * n + j];
            }
        }
        __syncthreads();
    }
    for (int e = tid; e < n * n; e += nt) {
        const int i = e / n, j = e - i * n;
        cinv[(size_t)j * n + i] = B[e];
    }
}

// ---- host side ---------------------------------------------------------------------------------------------
static inline int mg_grid(long long work) { return grid_for(work, 256, 148 * 8); }
// one resident wave of THIS kernel: a grid sized for 8 blocks per SM leaves a kernel that fits 5 or 6 (the reducing
// variant of the smoothing pass, the restriction) with a second, mostly empty wave -- measured +40 % on its run time
template <typename K>
static int mg_wave_grid(K kernel, long long work)
{
    // resident blocks of each kernel, by its address (kernels of one signature share this instantiation)
    static std::mutex mu;
    static std::unordered_map<const void *, int> known;
    int resident = 0;
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = known.find((const void *)kernel);
        if (it != known.end()) resident = it->second;
    }
    if (resident == 0) {
        int dev = 0, sms = 148, occ = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, 256, 0) != cudaSuccess || occ < 1) { cudaGetLastError(); occ = 4; }
        resident = sms * occ;
        std::lock_guard<std::mutex> lk(mu);
        known[(const void *)kernel] = resident;
    }
    return grid_for(work, 256, resident);
}
static inline VecView view(double *p, int stride, int off = 0) { VecView v; v.p = p; v.stride = stride; v.off = off; return v; }

bool mg_applicable(const GridDesc &g, int j0, int j1)
{
    if (g.type != RS_TYPE_2D && g.type != RS_TYPE_3D) return false;
    if (g.nspec < 1 || g.nspec > MG_MAXS) return false;
    const int d1 = j1 > j0 ? j1 - j0 : g.d[0];
    const long long cells = (long long)d1 * (g.ncell / g.d[0]);
    return cells >= MG_MIN_CELLS;
}

void mg_destroy(Mg *m) { delete m; }

// Lays out the hierarchy for the rank's part of the grid: planes [j0, j1) of the slowest dimension (j1 <= j0: the
// whole grid).  The grid is handled as a 3-D box with leading extents 1 (a 2-D grid (a, b) is (1, a, b)).
//
// Sharded solves.  A hierarchy built on the rank's slab alone (couplings across slab faces cut) is a block-Jacobi
// preconditioner over the ranks: measured at 2 GPUs, 256^3, 47 Krylov steps instead of 17 -- the cut is felt over the
// screening length of the operator.  So the finest levels are GLOBAL: same aggregates as on one GPU, every rank working
// on its planes of global-index level vectors, one halo exchange before each application of the level's operator (four
// per level and cycle).  A level can be global while every rank's planes are whole aggregates (even bounds); below
// that -- and below RS_MG_GLOBAL_LEVELS (default 3; at 2 GPUs, 256^3: 36 / 31 / 32 ms with 2 / 3 / 4) levels, where the planes per rank get few and an exchange costs
// more than the level's arithmetic -- the rank's coarse box is cut off and continues as a local hierarchy.
int mg_setup(Ctx &c, Mg **out, const GridDesc &g, int j0, int j1)
{
    std::unique_ptr<Mg> M(new Mg());
    const int nd = g.type == RS_TYPE_2D ? 2 : 3;
    int d[3] = {1, 1, 1}, cyc[3] = {0, 0, 0};
    for (int q = 0; q < nd; q++) { d[3 - nd + q] = g.d[q]; cyc[3 - nd + q] = g.cyc[q]; }
    const int slow = 3 - nd;   // position of the sharded (slowest) dimension in the box
    M->gstride = g.ncell;
    M->goff = 0;
    const bool sharded = j1 > j0 && j1 - j0 != g.d[0];
    const int nr = sharded && comm_active() ? comm_nranks() : 1, rank = nr > 1 ? comm_rank() : 0;
    // my planes at every global level; G = number of global levels
    int G = 0, myj0[MG_MAXLEV], myj1[MG_MAXLEV];
    if (nr > 1 && !getenv("RS_MG_LOCAL")) {
        const char *e = getenv("RS_MG_GLOBAL_LEVELS");
        const int gmax = e ? atoi(e) : 3;
        std::vector<int> lo(nr), hi(nr);
        for (int r = 0; r < nr; r++) rs_slab_bounds(g.d[0], nr, r, &lo[r], &hi[r]);
        int dl = g.d[0];
        while (G < gmax && G < MG_MAXLEV - 2) {
            bool ok = dl > 2;
            for (int r = 0; r < nr && ok; r++)
                ok = lo[r] % 2 == 0 && (hi[r] % 2 == 0 || hi[r] == dl) && hi[r] - lo[r] >= 2;
            if (!ok) break;
            myj0[G] = lo[rank]; myj1[G] = hi[rank];
            G++;
            for (int r = 0; r < nr; r++) { lo[r] /= 2; hi[r] = (hi[r] + 1) / 2; }
            dl = (dl + 1) / 2;
        }
        myj0[G] = lo[rank]; myj1[G] = hi[rank];   // my planes of the first local level
    }
    M->nglobal = G;
    if (sharded && G == 0) {   // purely slab-local
        M->goff = j0 * (g.ncell / g.d[0]);
        d[slow] = j1 - j0;
    }
    if (sharded) cyc[slow] = 0;   // a slab of a cyclic dimension is an open box: the wrap crosses ranks (not sharded today)
    int nlev = 0;
    for (;;) {
        MgLevel &L = M->lev[nlev];
        MgLevelDev &D = L.dev;
        if (G > 0 && nlev == G) d[slow] = myj1[G] - myj0[G];   // from here on: the rank's own box
        L.global = nlev < G;
        D.nspec = g.nspec;
        D.d1 = d[0]; D.d2 = d[1]; D.d3 = d[2];
        D.cyc1 = cyc[0]; D.cyc2 = cyc[1]; D.cyc3 = cyc[2];
        D.ncell = d[0] * d[1] * d[2];
        D.n = D.nspec * D.ncell;
        D.cb = 0; D.ce = D.ncell;
        if (L.global) {
            const int plane = D.ncell / d[slow];
            D.cb = myj0[nlev] * plane; D.ce = myj1[nlev] * plane;
            L.hp.nspec = D.nspec; L.hp.ncell = D.ncell; L.hp.plane = plane;
            L.hp.j0 = myj0[nlev]; L.hp.j1 = myj1[nlev];
            L.hp.has_lo = rank > 0; L.hp.has_hi = rank < nr - 1;
            if (nlev == G - 1) L.cjoff = myj0[G];
        }
        const size_t nslot = 7 + (size_t)D.nspec - 1;
        RS_TRY(L.ell.alloc(nslot * D.n));
        RS_TRY(L.dinv.alloc(D.nspec == 1 ? (size_t)D.n : (size_t)D.nspec * D.n));
        const bool lean = !L.global && D.nspec == 1;   // one-species local level: its passes share x and df only
        if (!lean) RS_TRY(L.r.alloc(D.n));
        if (!L.global) {   // global: below
            RS_TRY(L.x.alloc(D.n));
            if (!lean) RS_TRY(L.d.alloc(D.n));
            L.xp = L.x.p; L.dp = L.d.p;
        }
        if (lean) RS_TRY(L.df.alloc(D.n));
        if (nlev > 0) RS_TRY(L.b.alloc(D.n));
        D.ell = L.ell.p; D.dinv = L.dinv.p;
        nlev++;
        bool any = false;
        for (int q = 0; q < 3; q++) { L.f[q] = d[q] > 2 ? 2 : 1; any = any || L.f[q] == 2; }
        const bool more_global = nlev < G;   // a global level always hands on to the next one
        if (!more_global && nlev > G && (D.ncell <= MG_COARSEST_CELLS || !any || nlev == MG_MAXLEV)) { L.f[0] = L.f[1] = L.f[2] = 1; break; }
        for (int q = 0; q < 3; q++) d[q] = (d[q] + L.f[q] - 1) / L.f[q];
    }
    M->nlev = nlev;
    if (G > 0) {
        // x and d of the global levels: one peer-mapped arena when the halo planes can be pushed over NVLink
        size_t total = 0;
        for (int l = 0; l < G; l++) total += 2 * (((size_t)M->lev[l].dev.n + 31) & ~(size_t)31);
        double *mine = nullptr;
        void *alo = nullptr, *ahi = nullptr;
        M->push = !getenv("RS_NO_PUSH") && comm_pooled_arena(total, &mine, &alo, &ahi);
        size_t off = 0;
        for (int l = 0; l < G; l++) {
            MgLevel &L = M->lev[l];
            const size_t nl = ((size_t)L.dev.n + 31) & ~(size_t)31;
            if (M->push) {
                L.xp = mine + off; L.dp = mine + off + nl;
                L.x_lo = alo ? (double *)alo + off : nullptr; L.d_lo = alo ? (double *)alo + off + nl : nullptr;
                L.x_hi = ahi ? (double *)ahi + off : nullptr; L.d_hi = ahi ? (double *)ahi + off + nl : nullptr;
            } else {
                RS_TRY(L.x.alloc(L.dev.n)); RS_TRY(L.d.alloc(L.dev.n));
                RS_TRY(L.x.zero(c.stream)); RS_TRY(L.d.zero(c.stream));
                L.xp = L.x.p; L.dp = L.d.p;
            }
            off += 2 * nl;
        }
        RS_TRY(M->hticket.alloc(1));
        RS_TRY(M->hticket.zero(c.stream));
    }
    RS_TRY(M->lmax.alloc(nlev));
    for (int l = 0; l < nlev; l++) M->lev[l].dev.lmax = M->lmax.p + l;
    M->nc = M->lev[nlev - 1].dev.n;
    if (M->nc > 256 || nlev < 2) { set_error("multigrid: unusable hierarchy (%d levels, %d coarsest unknowns)", nlev, M->nc); return RS_ETOOBIG; }
    RS_TRY(M->cinv.alloc((size_t)M->nc * M->nc));
    RS_TRY(M->cwork.alloc(2 * (size_t)M->nc * M->nc));
    RS_TRY(M->flag.alloc(1));
    // the levels small enough for the one-launch tail (local levels only: a block cannot exchange halos)
    M->tail = nlev;
    for (int l = G; l < nlev; l++)
        if (M->lev[l].dev.ncell <= MG_TAIL_CELLS) { M->tail = l; break; }
    if (getenv("RS_NO_MG_TAIL")) M->tail = nlev;
    if (M->tail < nlev) {
        std::vector<MgTailLevel> tl(nlev);
        for (int l = 0; l < nlev; l++) {
            MgLevel &L = M->lev[l];
            tl[l].D = L.dev;
            tl[l].f1 = L.f[0]; tl[l].f2 = L.f[1]; tl[l].f3 = L.f[2];
            tl[l].b = L.b.p; tl[l].x = L.xp; tl[l].r = L.r.p; tl[l].d = L.dp; tl[l].df = L.df.p;
        }
        RS_TRY(M->tail_levels.alloc(sizeof(MgTailLevel) * nlev));
        RS_CUDA(cudaMemcpyAsync(M->tail_levels.p, tl.data(), sizeof(MgTailLevel) * nlev, cudaMemcpyHostToDevice, c.stream));
        RS_CUDA(cudaStreamSynchronize(c.stream));
    }
    *out = M.release();
    return RS_OK;
}

// Rebuilds every level from the matrix values the Krylov solver is about to use (once per Newton iteration).
int mg_update(Ctx &c, Mg &M, const int *rowptr, const int *col, const double *val, bool grid_order)
{
    cudaStream_t st = c.stream;
    RS_TRY(M.lmax.zero(st));
    RS_TRY(M.flag.zero(st));
    // one species: the kernel that forms a level's rows also leaves its smoother data (dinv, Gershgorin bound)
    const bool one = M.lev[0].dev.nspec == 1;
    bool smoother_done = false;   // of the level the loop is at
    {
        MgLevel &L = M.lev[0];
        const double nz = (double)L.dev.n * (6.0 + L.dev.nspec), rows = (double)(L.dev.ce - L.dev.cb);
        if (one && grid_order && M.goff == 0 && L.dev.ncell == M.gstride) {
            Launch La(c, K_PRECOND_SETUP, (4.0 + 56.0 + 28.0 + 8.0) * rows);
            mg_build_l0_grid_kernel<<<mg_wave_grid(mg_build_l0_grid_kernel, L.dev.ce - L.dev.cb), 256, 0, st>>>(L.dev, rowptr, val, L.ell.p, L.dinv.p,
                                                                                                           M.lmax.p);
            smoother_done = true;
        } else {
            RS_TRY(L.ell.zero(st));
            Launch La(c, K_PRECOND_SETUP, 12.0 * nz + 8.0 * nz);
            mg_build_l0_kernel<<<mg_grid((long long)L.dev.nspec * (L.dev.ce - L.dev.cb)), 256, 0, st>>>(L.dev, M.gstride, M.goff, rowptr, col, val, L.ell.p);
        }
    }
    for (int l = 0; l < M.nlev; l++) {
        MgLevel &L = M.lev[l];
        const double nz = (double)L.dev.n * (6.0 + L.dev.nspec);
        if (!smoother_done) {   // before the coarse build overwrites nothing of this level: order is free, kept as it was
            Launch La(c, K_PRECOND_SETUP, 8.0 * nz + 8.0 * L.dev.nspec * L.dev.n);
            mg_smoother_setup_kernel<<<grid_for(L.dev.ce - L.dev.cb, 128, 148 * 16), 128, 0, st>>>(L.dev, L.dinv.p, M.lmax.p + l);
        }
        smoother_done = false;
        if (l + 1 < M.nlev) {
            MgLevel &Cn = M.lev[l + 1];
            Launch La(c, K_PRECOND_SETUP, sizeof(mg_real) * nz + sizeof(mg_real) * nz / 8.0 + (one ? 8.0 * Cn.dev.n : 0.0));
            if (one) {
                mg_build_coarse1_kernel<<<mg_wave_grid(mg_build_coarse1_kernel, Cn.dev.ce - Cn.dev.cb), 256, 0, st>>>(L.dev, Cn.dev, L.f[0], L.f[1], L.f[2], L.cjoff, 0.5,
                                                                                                                 Cn.ell.p, Cn.dinv.p, M.lmax.p + l + 1);
                smoother_done = true;
            } else
                mg_build_coarse_kernel<<<mg_grid((long long)Cn.dev.nspec * (Cn.dev.ce - Cn.dev.cb)), 256, 0, st>>>(L.dev, Cn.dev, L.f[0], L.f[1], L.f[2], L.cjoff,
                                                                                                             0.5, Cn.ell.p);
        }
    }
    // a global level is one operator: every rank smooths with the same polynomial
    if (M.nglobal > 0) RS_TRY(comm_allreduce_max(c, M.lmax.p, M.nglobal));
    {
        Launch La(c, K_PRECOND_SETUP, 24.0 * M.nc * M.nc);
        mg_coarse_invert_kernel<<<1, 1024, 0, st>>>(M.lev[M.nlev - 1].dev, M.cwork.p, M.cinv.p, M.flag.p);
    }
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

// halo planes of x (or d) of a global level to the slab neighbours
static void mg_halo(Ctx &c, Mg &M, MgLevel &L, bool is_x)
{
    double *v = is_x ? L.xp : L.dp;
    if (!M.push) { comm_halo_exchange(c, v, L.hp); return; }
    const HaloSignal sig = comm_next_halo_signal();
    HaloPush hp;
    hp.lo = L.hp.has_lo ? (is_x ? L.x_lo : L.d_lo) : nullptr;
    hp.hi = L.hp.has_hi ? (is_x ? L.x_hi : L.d_hi) : nullptr;
    hp.plane = L.hp.plane;
    hp.signal = 1;
    hp.sig = sig;
    Launch La(c, K_HALO, 32.0 * L.hp.nspec * L.hp.plane);
    static const int push_blocks = getenv("RS_MG_PUSH_BLOCKS") ? atoi(getenv("RS_MG_PUSH_BLOCKS")) : 148;
    mg_halo_push_kernel<<<grid_for(2LL * L.hp.nspec * L.hp.plane, 256, push_blocks), 256, 0, c.stream>>>(L.hp.nspec, L.hp.ncell, L.hp.plane, L.hp.j0, L.hp.j1,
                                                                                            v, hp, M.hticket.p);
}

// resident blocks per SM the one-species level kernels are compiled for (RS_MG_MINB: 3..6, tuning)
static int mg_minb()
{
    static const int v = getenv("RS_MG_MINB") ? atoi(getenv("RS_MG_MINB")) : 4;
    return v;
}
#define MG_MINB_SWITCH(MB, ...)                                      \
    switch (mg_minb()) {                                             \
    case 3: { constexpr int MB = 3; __VA_ARGS__; } break;            \
    case 5: { constexpr int MB = 5; __VA_ARGS__; } break;            \
    case 6: { constexpr int MB = 6; __VA_ARGS__; } break;            \
    default: { constexpr int MB = 4; __VA_ARGS__; } break;           \
    }

template <int NS>
static void mg_cycle_level(Ctx &c, Mg &M, int l, VecView b, VecView xout, double *partials, unsigned *ticket, double *dot_out,
                           const PeerRed &pr)
{
    cudaStream_t st = c.stream;
    MgLevel &L = M.lev[l];
    const MgLevelDev &D = L.dev;
    const double n = (double)D.nspec * (D.ce - D.cb), mat = (double)sizeof(mg_real) * n * (6.0 + D.nspec);
    // statistics: the smoothing passes and the transfers of the finest level on their own, every level below as one class
    const int k_smooth = l == 0 ? K_MG_SMOOTH_L0 : K_MG_COARSE, k_transfer = l == 0 ? K_MG_TRANSFER_L0 : K_MG_COARSE;
    if (l == M.tail && !(l == 0 && dot_out && pr.nranks > 1)) {
        // this level and everything below it in one block
        double bytes = 0.0;
        for (int q = l; q < M.nlev; q++) bytes += 4.0 * sizeof(mg_real) * M.lev[q].dev.n * (6.0 + D.nspec) + 150.0 * M.lev[q].dev.n;
        Launch La(c, K_MG_COARSE, bytes);
        const VecView none0 = view(nullptr, 0);
        mg_tail_kernel<NS><<<1, MG_TAIL_THREADS, 0, st>>>(reinterpret_cast<const MgTailLevel *>(M.tail_levels.p), l, M.nlev, M.nc, M.cinv.p,
                                                         l == 0 ? b : none0, l == 0 ? xout : none0, dot_out ? 1 : 0, dot_out);
        return;
    }
    if (l == M.nlev - 1) {
        Launch La(c, K_MG_COARSE, 8.0 * M.nc * M.nc + 16.0 * M.nc);
        mg_coarse_solve_kernel<<<1, 256, sizeof(double) * M.nc, st>>>(M.nc, M.cinv.p, b.p, xout.p);
        return;
    }
    MgLevel &Cn = M.lev[l + 1];
    const VecView x = view(L.xp, D.ncell), r = view(L.r.p, D.ncell), d = view(L.dp, D.ncell);
    const VecView none = view(nullptr, 0);
    if (NS == 1 && !L.global) {   // pre-smoothing from a zero guess: two Chebyshev steps, one pass, d never stored
        Launch La(c, k_smooth, mat + (sizeof(mg_real) + 16.0) * n);   // + dinv, b, x out
        MG_MINB_SWITCH(MB, mg_pre_kernel<MB><<<mg_wave_grid(mg_pre_kernel<MB>, D.ce - D.cb), 256, 0, st>>>(D, b, x));
    } else {
        {
            Launch La(c, k_smooth, 16.0 * n + 8.0 * D.nspec * n);
            mg_first_kernel<NS><<<mg_wave_grid(mg_first_kernel<NS>, D.ce - D.cb), 256, 0, st>>>(D, b, d);
        }
        if (L.global) mg_halo(c, M, L, false);
        Launch La(c, k_smooth, mat + 24.0 * n + 8.0 * D.nspec * n);
        mg_second_kernel<NS, false><<<mg_wave_grid(mg_second_kernel<NS, false>, D.ce - D.cb), 256, 0, st>>>(D, b, d, none, x, none, nullptr, nullptr,
                                                                                                         nullptr, PeerRed());
    }
    if (L.global) mg_halo(c, M, L, true);
    {   // residual, summed over the aggregates
        Launch La(c, k_transfer, mat + 16.0 * n + 8.0 * Cn.dev.n);
        const VecView bcv = view(Cn.b.p, Cn.dev.ncell);
        if (NS == 1) {
            MG_MINB_SWITCH(MB, mg_resid_restrict_kernel<1, MB><<<mg_wave_grid(mg_resid_restrict_kernel<1, MB>, Cn.dev.ce - Cn.dev.cb), 256, 0, st>>>(
                                   D, L.f[0], L.f[1], L.f[2], Cn.dev, L.cjoff, b, x, bcv));
        } else
            mg_resid_restrict_kernel<NS, 1><<<mg_wave_grid(mg_resid_restrict_kernel<NS, 1>, Cn.dev.ce - Cn.dev.cb), 256, 0, st>>>(D, L.f[0], L.f[1], L.f[2], Cn.dev,
                                                                                                                   L.cjoff, b, x, bcv);
    }
    // deepest global level: acknowledge that the x halo has been read (see mg_halo_raise_kernel)
    const bool ack = L.global && l == M.nglobal - 1 && M.push;
    if (ack) {
        Launch La(c, K_HALO, 16.0);
        mg_halo_raise_kernel<<<1, 1, 0, st>>>(comm_next_halo_signal());
    }
    mg_cycle_level<NS>(c, M, l + 1, view(Cn.b.p, Cn.dev.ncell), view(Cn.xp, Cn.dev.ncell), nullptr, nullptr, nullptr, PeerRed());
    if (NS == 1 && !L.global) {   // prolongation folded into the post-smoothing; its two passes share only d (single precision)
        const VecView xc = view(Cn.xp, Cn.dev.ncell);
        const double ms = sizeof(mg_real);
        {
            Launch La(c, k_smooth, mat + (ms + 16.0 + ms) * n + 8.0 * Cn.dev.n);   // + dinv, b, x, d out; xc
            MG_MINB_SWITCH(MB, mg_post_first_kernel<MB><<<mg_wave_grid(mg_post_first_kernel<MB>, D.ce - D.cb), 256, 0, st>>>(
                                   D, L.f[0], L.f[1], L.f[2], Cn.dev.d2, Cn.dev.d3, b, x, xc, L.df.p));
        }
        Launch La(c, k_smooth, mat + (ms + ms + 16.0 + (dot_out ? 8.0 : 0.0)) * n + 8.0 * Cn.dev.n);   // + dinv, d, x, xout[, b]; xc
        if (dot_out)
            mg_post_second_kernel<true, 3><<<mg_wave_grid(mg_post_second_kernel<true, 3>, D.ce - D.cb), 256, 0, st>>>(
                D, L.f[0], L.f[1], L.f[2], Cn.dev.d2, Cn.dev.d3, x, xc, L.df.p, xout, b, partials, ticket, dot_out, pr);
        else {
            MG_MINB_SWITCH(MB, mg_post_second_kernel<false, MB><<<mg_wave_grid(mg_post_second_kernel<false, MB>, D.ce - D.cb), 256, 0, st>>>(
                                   D, L.f[0], L.f[1], L.f[2], Cn.dev.d2, Cn.dev.d3, x, xc, L.df.p, xout, none, nullptr, nullptr, nullptr, PeerRed()));
        }
        return;
    }
    {
        Launch La(c, k_transfer, 16.0 * n + 8.0 * Cn.dev.n);
        mg_prolong_kernel<<<mg_wave_grid(mg_prolong_kernel, D.ce - D.cb), 256, 0, st>>>(D, L.f[0], L.f[1], L.f[2], Cn.dev.d2, Cn.dev.d3, L.cjoff, view(Cn.xp, Cn.dev.ncell), x);
    }
    if (ack) {   // the neighbours have finished reading the x planes the push below overwrites
        Launch La(c, K_HALO, 16.0);
        mg_halo_wait_kernel<<<1, 32, 0, st>>>(comm_next_halo_signal());
    }
    if (L.global) mg_halo(c, M, L, true);
    {   // post-smoothing: the same two Chebyshev steps, two passes
        Launch La(c, k_smooth, mat + 32.0 * n + 8.0 * D.nspec * n);
        mg_resid_first_kernel<NS><<<mg_wave_grid(mg_resid_first_kernel<NS>, D.ce - D.cb), 256, 0, st>>>(D, b, x, r, d);
    }
    if (L.global) mg_halo(c, M, L, false);
    {
        Launch La(c, k_smooth, mat + 32.0 * n + 8.0 * D.nspec * n);
        if (dot_out)
            mg_second_kernel<NS, true><<<mg_wave_grid(mg_second_kernel<NS, true>, D.ce - D.cb), 256, 0, st>>>(D, r, d, x, xout, b, partials, ticket,
                                                                                                            dot_out, pr);
        else
            mg_second_kernel<NS, false><<<mg_wave_grid(mg_second_kernel<NS, false>, D.ce - D.cb), 256, 0, st>>>(D, r, d, x, xout, none, nullptr, nullptr,
                                                                                                             nullptr, PeerRed());
    }
}

// z = M^-1 r: one V-cycle.  r and z are Krylov vectors in the solver's global indexing (only the rank's slab is
// read and written); dot_out != nullptr: *dot_out = <r, z> over the slab, all-reduced over the ranks (pr).
int mg_apply(Ctx &c, Mg &M, const double *r, double *z, double *partials, unsigned *ticket, double *dot_out, const PeerRed &pr)
{
    // level 0 addresses the Krylov vectors directly: cells of the rank's box (slab-local hierarchy: offset to its first
    // plane) or of the whole grid (global top level: the level's own cell range picks the rank's planes)
    const VecView b = view(const_cast<double *>(r), M.gstride, M.goff), xo = view(z, M.gstride, M.goff);
    const int ns = M.lev[0].dev.nspec;
    if (ns == 1) mg_cycle_level<1>(c, M, 0, b, xo, partials, ticket, dot_out, pr);
    else if (ns == 2) mg_cycle_level<2>(c, M, 0, b, xo, partials, ticket, dot_out, pr);
    else mg_cycle_level<0>(c, M, 0, b, xo, partials, ticket, dot_out, pr);
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

int mg_levels(const Mg &M) { return M.nlev; }

// out2 (device) = { max |a_ij - a_ji| , max |a_ij| } of the matrix the hierarchy was last built from (mg_update first)
int mg_asymmetry(Ctx &c, Mg &M, double *partials, unsigned *ticket, double *out2)
{
    const MgLevelDev &D = M.lev[0].dev;
    Launch La(c, K_PRECOND_SETUP, 2.0 * sizeof(mg_real) * (double)D.nspec * (D.ce - D.cb) * (6.0 + D.nspec));
    mg_asymmetry_kernel<<<mg_wave_grid(mg_asymmetry_kernel, D.ce - D.cb), 256, 0, c.stream>>>(D, partials, ticket, out2);
    RS_CUDA(cudaGetLastError());
    return RS_OK;
}

}  // namespace rs
