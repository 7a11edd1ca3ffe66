"""Prototype (scipy) of the aggregation-multigrid preconditioner: iteration counts of PCG for the variants
considered in DESIGN.md before any CUDA was written.  Piecewise-constant 2x2x2 aggregation, Galerkin coarse
operators (optionally with the off-diagonal rescaling that restores the h^-2 scaling of the diffusive part),
damped-Jacobi / Chebyshev smoothing, V-cycle, optional slab-local (block-diagonal over slabs) variant."""
import sys, time
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


def lap1d(n, dirichlet=True, cyclic=False):
    e = np.ones(n)
    T = sp.diags([e[:-1], -2 * e, e[:-1]], [-1, 0, 1], format="lil")
    if not dirichlet:
        T[0, 0] = -1; T[-1, -1] = -1
    if cyclic:
        T[0, -1] = 1; T[-1, 0] = 1
    return T.tocsr()


def grid_matrix(dims, r, dirichlet=True, rvar=None):
    """-(Laplacian - r) on a box grid, dims[0] slowest.  Positive definite."""
    mats = []
    nd = len(dims)
    for q in range(nd):
        ops = [sp.identity(d, format="csr") for d in dims]
        ops[q] = lap1d(dims[q], dirichlet)
        M = ops[0]
        for o in ops[1:]:
            M = sp.kron(M, o, format="csr")
        mats.append(M)
    A = -sum(mats)
    n = A.shape[0]
    rr = np.full(n, r) if rvar is None else rvar
    return (A + sp.diags(rr)).tocsr()


def agg_P(dims):
    """piecewise-constant prolongation for 2x aggregation in every dim with extent > 2"""
    cd = [(d + 1) // 2 if d > 2 else d for d in dims]
    idx = np.indices(dims).reshape(len(dims), -1)
    cidx = [idx[q] // 2 if dims[q] > 2 else idx[q] for q in range(len(dims))]
    cflat = np.ravel_multi_index(cidx, cd)
    n = int(np.prod(dims)); nc = int(np.prod(cd))
    P = sp.csr_matrix((np.ones(n), (np.arange(n), cflat)), shape=(n, nc))
    return P, cd


def galerkin(A, P, fix):
    Ac = (P.T @ A @ P).tocsr()
    if fix:
        # halve the off-diagonals, keep the row sums (the zeroth-order part is summed exactly)
        D = Ac.diagonal()
        off = Ac - sp.diags(D)
        rowsum = np.asarray(Ac.sum(axis=1)).ravel()
        off = off * fix
        Ac = (off + sp.diags(rowsum - np.asarray(off.sum(axis=1)).ravel())).tocsr()
    return Ac


class MG:
    def __init__(self, A, dims, fix=0.0, nu=1, omega=0.8, cheb=0, min_cells=64, slabs=1, wcycle=False):
        self.levels = []
        self.nu, self.omega, self.cheb, self.w = nu, omega, cheb, wcycle
        if slabs > 1:   # slab-local: drop couplings across slab faces (slowest dim)
            n = A.shape[0]
            plane = n // dims[0]
            sl = (np.arange(n) // plane) * slabs // dims[0]
            Ac = A.tocoo()
            keep = sl[Ac.row] == sl[Ac.col]
            A = sp.csr_matrix((Ac.data[keep], (Ac.row[keep], Ac.col[keep])), shape=A.shape)
            # aggregation never crosses slabs if slab bounds are multiples of 2^levels: assume so
        while True:
            lev = dict(A=A, dims=list(dims), dinv=1.0 / A.diagonal())
            if cheb:
                # eigenvalue bound of D^-1 A by Gershgorin-ish: use power iteration estimate
                lev["lmax"] = self.est_lmax(A, lev["dinv"])
            self.levels.append(lev)
            if int(np.prod(dims)) <= min_cells or all(d <= 2 for d in dims):
                break
            P, cd = agg_P(dims)
            lev["P"] = P
            A = galerkin(A, P, fix)
            dims = cd
        Ad = self.levels[-1]["A"].toarray()
        self.coarse_inv = np.linalg.inv(Ad)

    @staticmethod
    def est_lmax(A, dinv):
        # Gershgorin bound on D^-1 A
        absA = abs(A)
        return float((np.asarray(absA.sum(axis=1)).ravel() * abs(dinv)).max())

    def smooth(self, lev, x, b, first_zero):
        A, dinv = lev["A"], lev["dinv"]
        if self.cheb:
            lmax = lev["lmax"]; lmin = lmax / 8.0   # smooth the upper part of the spectrum
            theta = 0.5 * (lmax + lmin); delta = 0.5 * (lmax - lmin)
            sigma = theta / delta; rho = 1.0 / sigma
            r = b - A @ x if not first_zero else b.copy()
            d = dinv * r / theta
            x = x + d
            for _ in range(self.cheb - 1):
                r = r - A @ d
                rho_new = 1.0 / (2 * sigma - rho)
                d = rho_new * rho * d + 2 * rho_new / delta * (dinv * r)
                rho = rho_new
                x = x + d
            return x
        for s in range(self.nu):
            if first_zero and s == 0:
                x = self.omega * dinv * b
            else:
                x = x + self.omega * dinv * (b - A @ x)
        return x

    def cycle(self, l, b):
        lev = self.levels[l]
        if l == len(self.levels) - 1:
            return self.coarse_inv @ b
        x = self.smooth(lev, np.zeros_like(b), b, True)
        r = b - lev["A"] @ x
        bc = lev["P"].T @ r
        ec = self.cycle(l + 1, bc)
        if self.w and l + 1 < len(self.levels) - 1:
            # second visit: residual correction on the coarse level
            Ac = self.levels[l + 1]["A"]
            ec = ec + self.cycle(l + 1, bc - Ac @ ec)
        x = x + lev["P"] @ ec
        x = self.smooth(lev, x, b, False)
        return x

    def __call__(self, b):
        return self.cycle(0, b)

    def cost(self):
        """matrix passes per cycle in units of the fine matrix"""
        nn = [lv["A"].nnz for lv in self.levels]
        per = (2 * self.nu if not self.cheb else 2 * self.cheb)  # first pre-smoothing step is free, + residual
        tot = 0.0
        mult = 1
        for l, z in enumerate(nn[:-1]):
            tot += mult * per * z / nn[0]
            if self.w and l >= 0:
                mult *= 2 if l + 1 < len(nn) - 1 else 1
        return tot


def pcg(A, b, M, rtol=1e-8, maxit=2000):
    x = np.zeros_like(b); r = b.copy(); z = M(r); p = z.copy(); rz = r @ z
    bn = np.linalg.norm(b)
    for it in range(1, maxit + 1):
        q = A @ p
        a = rz / (p @ q)
        x += a * p; r -= a * q
        if np.linalg.norm(r) <= rtol * bn:
            return x, it
        z = M(r); rzn = r @ z
        p = z + (rzn / rz) * p; rz = rzn
    return x, maxit


def run(name, A, dims, b=None, **kw):
    n = A.shape[0]
    if b is None:
        rng = np.random.default_rng(1)
        b = rng.standard_normal(n)
    t0 = time.time()
    M = MG(A, dims, **kw)
    x, it = pcg(A, b, M)
    dinv = 1.0 / A.diagonal()
    res = np.linalg.norm(b - A @ x) / np.linalg.norm(b)
    print(f"{name:28s} {str(kw):70s} levels={len(M.levels)} its={it:4d} passes/cycle={M.cost():.2f} "
          f"total_passes={(M.cost() + 1) * it:.0f} res={res:.1e} ({time.time() - t0:.1f}s)", flush=True)
    return it


if __name__ == "__main__":
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    cases = {
        "C5 r=.025 dirichlet": (grid_matrix([S, S, S], 0.025), [S, S, S]),
        "3D r=1e-4 zero-flux": (grid_matrix([S, S, S], 1e-4, dirichlet=False), [S, S, S]),
    }
    for name, (A, dims) in cases.items():
        n = A.shape[0]
        rng = np.random.default_rng(1)
        b = rng.standard_normal(n)
        x, itj = pcg(A, b, lambda r, d=1.0 / A.diagonal(): d * r)
        print(f"{name}: Jacobi-CG its={itj}")
        for kw in [dict(fix=0.0, nu=1, omega=0.8), dict(fix=0.5, nu=1, omega=0.8), dict(fix=0.0, nu=2, omega=0.8),
                   dict(fix=0.5, nu=2, omega=0.8), dict(fix=0.0, cheb=2), dict(fix=0.5, cheb=2), dict(fix=0.5, cheb=3),
                   dict(fix=0.0, nu=1, omega=0.8, wcycle=True), dict(fix=0.5, nu=1, omega=0.8, slabs=8),
                   dict(fix=0.0, nu=1, omega=0.8, slabs=8)]:
            run(name, A, dims, b, **kw)
