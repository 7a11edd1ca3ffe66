"""The block-tridiagonal preconditioner (rootsolve_b200/csrc/btd.cu): on the structure of sparse1d
(src/sparsecoeff.f:5-83) it is an exact factorisation -- the role of nsfc/nnfc/nntc (src/sparse.f:1937-2563) -- so the
Krylov solve around it must finish in a step or two and agree with scipy's direct solve; on a cyclic grid and on a general
structure with a few entries off the band a handful of steps more."""
import ctypes as C

import numpy as np
import pytest
import scipy.sparse.linalg as spla

import helpers as H
from test_multigrid import grid_matrix

pytestmark = pytest.mark.gpu

BTD = 4


def solve_1d(lib, ian, jan, val, b, nspec, nb, cyc, precond=4, max_krylov=200, lin_rtol=1e-12):
    from rootsolve_b200 import _capi
    n = len(b)
    kc = _capi.KrylovControl(precond=precond, max_krylov=max_krylov, lin_rtol=lin_rtol)
    x = np.zeros(n)
    it, rr = C.c_int(0), C.c_double(0.0)
    used = np.zeros(2, dtype=np.int32)
    h = np.asarray([2, nspec, nb, 1, 1, cyc, 0, 0], dtype=np.int32)
    _capi.check(lib.rs_linsolve_grid(n, H.ip(ian), H.ip(jan), H.dp(val), H.dp(b), H.ip(h), C.byref(kc), H.dp(x), C.byref(it),
                                     C.byref(rr), H.ip(used)))
    return x, it.value, rr.value, used


CASES_1D = [(1, 2, 0), (1, 3, 0), (1, 7, 0), (1, 64, 0), (1, 1000, 0), (1, 4097, 0), (2, 33, 0), (3, 500, 0), (5, 1000, 0),
            (8, 129, 0), (5, 20000, 0), (1, 50, 1), (3, 77, 1), (5, 1024, 1)]


@pytest.mark.parametrize("nspec,nb,cyc", CASES_1D, ids=str)
def test_btd_is_exact_on_1d_grids(built, oracle, nspec, nb, cyc):
    rng = np.random.default_rng(11)
    # diffusion-dominated (tiny zeroth-order term), lopsided faces (advection), species coupled inside a box
    ian, jan, val, A = grid_matrix(oracle, "1D", nspec, [nb], [cyc], rng, react=1e-4, advect=0.3, couple=0.3 if nspec > 1 else 0.0)
    n = A.shape[0]
    xs = rng.standard_normal(n)
    b = A @ xs
    for precond in (BTD, -1):       # asked for, and what AUTO picks on a 1-D grid
        x, it, rr, used = solve_1d(built, ian, jan, val, b, nspec, nb, cyc, precond)
        assert used[0] == BTD, used
        assert rr <= 1e-10, rr
        assert it <= (3 if not cyc else 16), it
        xd = spla.spsolve(A.tocsc(), b)
        assert np.max(np.abs(x - xd)) <= 1e-7 * np.max(np.abs(xd))


def test_btd_beats_jacobi_on_a_bare_laplacian(built, oracle):
    """1000 boxes, no zeroth-order term to speak of: condition number ~ 4e5.  Jacobi-preconditioned Krylov needs
    thousands of steps (or stalls); the band solve one."""
    rng = np.random.default_rng(5)
    ian, jan, val, A = grid_matrix(oracle, "1D", 1, [1000], [0], rng, react=1e-6)
    b = rng.standard_normal(1000)
    # condition number 4e6: a backward-stable solve leaves a relative residual of about eps * cond
    x, it, rr, used = solve_1d(built, ian, jan, val, b, 1, 1000, 0, BTD, lin_rtol=1e-8)
    assert used[0] == BTD and it <= 2 and rr <= 1e-8, (used, it, rr)
    xd = spla.spsolve(A.tocsc(), b)
    assert np.max(np.abs(x - xd)) <= 1e-6 * np.max(np.abs(xd))


def test_whole_solves_pick_the_band_solve(built, oracle, tmp_path):
    """steady.1D (config C2's model), the numerically detected structure of man/stodes.Rd's sediment model (tridiagonal
    plus five entries far from the diagonal) and the user model started where its Jacobian is a bare Laplacian"""
    import rootsolve_b200 as rs
    from test_host_api import build_user_model
    nb, ns = 3000, 5
    out = rs.steady_1D(np.full(ns * nb, 1.0), nspec=ns, method="stodes", func="rt1d", parms=H.P_RT1D, positive=True,
                       ipar=[ns, nb, 1, 1, 0], **H.TOL)
    A = out.attributes
    assert A["steady"] and A["precond"] == "btd" and A["krylov_iters"] <= 3 * len(A["precis"]), A["krylov_iters"]
    ref = oracle.stsparse(np.full(ns * nb, 1.0), oracle.model("rt1d", H.P_RT1D), sparsetype="1D", nnz=[ns, nb, 0],
                          ipar=[ns, nb, 1, 1, 0], lrw=4000000, positive=True, **H.TOL)
    ok, worst = H.within_tol(out["y"], ref["y"])
    assert ok, worst
    y0 = np.random.default_rng(20260101).uniform(size=1000)
    out = rs.stodes(y0, func="sediment", parms=H.P_SEDIMENT, positive=True, ipar=[1, 1000])
    A = out.attributes
    assert A["steady"] and A["precond"] == "btd" and A["krylov_iters"] <= 12 * len(A["precis"]), A["krylov_iters"]
    so = build_user_model(tmp_path)
    out = rs.steady_1D(np.full(500, 5.0), func="fisher1d", dllname=so, parms=[1.0, 0.2, 0.5, 10.0, 1.0], nspec=1, positive=True,
                       method="stodes", **H.TOL)
    A = out.attributes
    assert A["steady"] and A["precond"] == "btd" and A["krylov_iters"] <= 3 * len(A["precis"]), A["krylov_iters"]
