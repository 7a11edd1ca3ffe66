"""The reference's own examples of the stodes path that live in the diffusion-dominated / nonsymmetric regime,
against the oracle ("what's missing" 2):
  * vignettes/rootSolve.Rnw:615-663 -- steady.2D, 100 x 100, zero-flux boundaries, quadratic consumption, ten
    producing cells; the only example of this path with a published run time;
  * man/stodes.Rd:225-284 -- 1000 sediment layers, Kz/dx^2 = 1e4 against decay = 5, non-local exchange, the default
    sparsetype = "sparseint", and the sparsereturn -> sparsejan round trip."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def test_vignette_2d_zero_flux_quadratic(built, oracle):
    import rootsolve_b200 as rs
    n = 100
    ipar = H.diff2dq_ipar(n)
    y0 = np.full((n, n), 10.0)
    ref = oracle.stsparse(y0.ravel(), oracle.model("diff2dq", H.P_DIFF2DQ), sparsetype="2D", nnz=[1, n, n, 0, 0], ipar=ipar,
                          lrw=1000000, positive=True, **H.TOL)
    # called as the vignette calls it: steady.2D(Conc0, func, parms = NULL, pos = TRUE, dimens = c(n, n), lrw = 1e6, tol 1e-10)
    out = rs.steady_2D(y0, func="diff2dq", parms=H.P_DIFF2DQ, dimens=[n, n], positive=True, lrw=1000000, ipar=ipar,
                       return_groups=True, **H.TOL)
    A = out.attributes
    assert ref["steady"] and A["steady"]
    assert A["dims"]["nnz"] == ref["dims"]["nnz"] == 49600 and A["dims"]["ngp"] == ref["dims"]["ngp"] == 7
    assert np.array_equal(A["igp"], ref["igp"]) and np.array_equal(A["jgp"], ref["jgp"])
    assert len(A["precis"]) == len(ref["precis"])
    assert abs(A["precis"][0] - ref["precis"][0]) <= 1e-12 * ref["precis"][0]
    ok, worst = H.within_tol(out["y"], ref["y"])
    assert ok, worst


def _sediment(n=1000, seed=20260101):
    return np.random.default_rng(seed).uniform(size=n), [1, n]


def test_man_stodes_sediment_sparseint(built, oracle):
    """stodes(runif(N), func = model, parms = NULL, positive = TRUE, decay = 5, ing = 20): default tolerances,
    structure detected numerically"""
    import rootsolve_b200 as rs
    y0, ipar = _sediment()
    ref = oracle.stsparse(y0, oracle.model("sediment", H.P_SEDIMENT), sparsetype="sparseint", nnz=None, ipar=ipar, positive=True)
    out = rs.stodes(y0, func="sediment", parms=H.P_SEDIMENT, positive=True, ipar=ipar, return_groups=True)
    A = out.attributes
    assert ref["steady"] and A["steady"]
    assert A["dims"]["nnz"] == ref["dims"]["nnz"] == 3003 and A["dims"]["ngp"] == ref["dims"]["ngp"]
    assert np.array_equal(A["igp"], ref["igp"]) and np.array_equal(A["jgp"], ref["jgp"])
    ok, worst = H.within_tol(out["y"], ref["y"], atol=1e-8, rtol=1e-6)
    assert ok, worst


def test_man_stodes_sparsereturn_sparsejan_roundtrip(built, oracle):
    """man/stodes.Rd:264-284: the structure returned by sparsetype = "sparsereturn" is fed back as "sparsejan",
    through stodes and through steady.1D(jactype = "sparsejan")"""
    import rootsolve_b200 as rs
    y0, ipar = _sediment()
    ref = oracle.stsparse(y0, oracle.model("sediment", H.P_SEDIMENT), sparsetype="sparseint", nnz=None, ipar=ipar, positive=True)
    sp = rs.stodes(y0, func="sediment", parms=H.P_SEDIMENT, sparsetype="sparsereturn", ipar=ipar)
    rs_ = oracle.stsparse(y0, oracle.model("sediment", H.P_SEDIMENT), sparsetype="sparsereturn", nnz=None, ipar=ipar)
    assert np.array_equal(sp.attributes["ian"], rs_["ian"]) and np.array_equal(sp.attributes["jan"], rs_["jan"])
    inz = np.concatenate([sp.attributes["ian"], sp.attributes["jan"]])
    s2 = rs.stodes(y0, func="sediment", parms=H.P_SEDIMENT, positive=True, sparsetype="sparsejan", inz=inz, ipar=ipar)
    s2b = rs.steady_1D(y0, func="sediment", parms=H.P_SEDIMENT, method="stodes", nspec=1, jactype="sparsejan", inz=inz,
                       ipar=ipar, positive=True)
    for o in (s2, s2b):
        assert o.attributes["steady"]
        ok, worst = H.within_tol(o["y"], ref["y"], atol=1e-8, rtol=1e-6)
        assert ok, worst
