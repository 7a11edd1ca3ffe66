"""Shared workload definitions for the tests (BASELINE.json configs at test sizes)."""
import ctypes as C

import numpy as np

TOL = dict(atol=1e-10, rtol=1e-10, ctol=1e-10)

# parameter vectors of the bundled models (rootsolve_b200/models/model_math.h)
P_DIFF2D_LIN = [1.0, 1.0, 1.0, 1.0, 0.025, 1.0, 1.0]
P_DIFF2D_MONOD = [1.0, 1.0, 1.0, 1.0, 0.25, 1.0, 1.0]
P_DIFF3D = [1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 0.025, 1.0]
# D1 D2 dx dy dz k m BNDu BNDv -- m > k*BNDu keeps dv/dv < 0 everywhere (a decaying, not a growing, species v)
P_RD3D = [1.0, 0.5, 1.0, 1.0, 1.0, 0.01, 0.2, 10.0, 1.0, 0.0]
P_RT1D = [1.0, 1.0, 1.0, 0.5, 0.5, 1.0, 0.1]
# man/steady.2D.Rd:193-196 + the two constants 10 of diffusion2Db: Dx Dy dx dy r BND extra exchange
P_DIFF2DB = [1.0, 1.0, 1.0, 1.0, 0.025, 1.0, 10.0, 10.0]
# man/steady.1D.Rd:137-145: v dx r p O2sat O2_0 BOD_0 ksO2
P_O2BOD = [1e2, 100.0, 0.1, 0.1, 300.0, 50.0, 1500.0, 10.0]


def ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def make_map(mask, nspec=1):
    """The presence map of sparsetype '2Dmap'/'3Dmap' (src/testMap.f): mask is a boolean array in stodes' dimension
    order (last index fastest), the same for every species.  Returns (pres, inv): pres[M-1] = new number of
    full-grid state M (0 if absent), inv[c] = full-grid number of unknown c (both 1-based values)."""
    m = np.tile(np.asarray(mask, dtype=bool).ravel(), nspec)
    pres = np.zeros(m.size, dtype=np.int32)
    pres[m] = np.arange(1, int(m.sum()) + 1, dtype=np.int32)
    inv = (np.nonzero(m)[0] + 1).astype(np.int32)
    return pres, inv


def ball_mask(dims, shrink=0.5):
    """cells whose centre lies inside the ellipsoid inscribed in the grid (a lake, a sphere...)"""
    axes = [(np.arange(d) + 0.5) / d - 0.5 for d in dims]
    g = np.meshgrid(*axes, indexing="ij")
    return sum(a * a for a in g) <= shrink * shrink


def diffmap_ipar(dims3, pres, inv):
    """ipar of the bundled `diffmap` model (model_math.h): nspec, d1, d2, d3, unused, n, inv[n], pres[ncell]"""
    return np.concatenate([[1, dims3[0], dims3[1], dims3[2], 0, len(inv)], inv, pres]).astype(np.int32)


def within_tol(y, yref, atol=1e-10, rtol=1e-10, factor=1.0):
    """The parity bound of BASELINE.json: |y - yref| <= atol + rtol*|yref| (times factor)."""
    err = np.abs(np.asarray(y) - np.asarray(yref))
    bound = factor * (atol + rtol * np.abs(yref))
    return bool(np.all(err <= bound)), float(np.max(err / bound))


def ms_setup(d, a=3.0, r=0.025, D=1.0, h=1.0, bnd=1.0):
    """Manufactured solution of the bundled `diff3dms` model (model_math.h): parameters, the rpar sine tables and
    the closed-form discrete steady state y* = bnd + a * sj x sk x sl on the grid d = (d1, d2, d3)."""
    tabs = [np.sin(np.pi * (np.arange(q) + 1) / (q + 1)) for q in d]
    lam = sum(D / h ** 2 * 2.0 * (1.0 - np.cos(np.pi / (q + 1))) for q in d)
    parms = [D, D, D, h, h, h, r, bnd, r * bnd, a * (lam + r)]
    ystar = bnd + a * np.einsum("j,k,l->jkl", *tabs).ravel()
    return parms, np.concatenate(tabs), ystar


# vignettes/rootSolve.Rnw:635-643: Dx Dy dx dy r p ; the ten producing cells are drawn by runif in the vignette,
# here from default_rng(20260101) (seeds)
P_DIFF2DQ = [1.5, 1.5, 1.0, 1.0, 0.01, 20.0]


def diff2dq_ipar(n, nsrc=10, seed=20260101):
    rng = np.random.default_rng(seed)
    cells = np.unique(np.stack([rng.integers(0, n, nsrc), rng.integers(0, n, nsrc)], 1), axis=0)
    src = (cells[:, 1] * n + cells[:, 0]).astype(np.int32)
    return [1, n, n, 1, len(src)] + [int(v) for v in src]


# man/stodes.Rd:243-251: v flux dx Kz decay ing
P_SEDIMENT = [0.1, 10.0, 0.01, 1.0, 5.0, 20.0]


def host_residual(oracle, model, parms, ipar, y, rpar=None):
    """f(y) evaluated by the oracle's host model (independent of every CUDA kernel)"""
    func = C.cast(oracle.model(model, parms), oracle.DERIV)
    n = len(y)
    f = np.zeros(n)
    ipfull = np.concatenate([[0, 0 if rpar is None else len(rpar), 3 + len(ipar)], ipar]).astype(np.int32)
    out = np.zeros(1) if rpar is None else np.ascontiguousarray(rpar, dtype=np.float64)
    neq, t = C.c_int(n), C.c_double(0.0)
    func(C.byref(neq), C.byref(t), dp(np.ascontiguousarray(y)), dp(f), dp(out), ip(ipfull))
    return f


def accept(y, f, atol=1e-10, rtol=1e-10):
    """the reference's own acceptance test (src/sparse.f:126-129): max |f / ewt| with ewt = rtol|y| + atol"""
    return float(np.max(np.abs(f) / (rtol * np.abs(y) + atol)))
