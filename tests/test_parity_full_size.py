"""Parity at the sizes that are benchmarked ("parity protocol where the reference cannot run"):
(1) oracle-vs-GPU state parity at the largest sizes the oracle's Yale LDU does in seconds to a minute,
(2) at BASELINE config C5's full size (steady.3D 256^3, the bench workload): structure and column grouping
    bit-exact against the oracle's generators (no LDU involved), and the reference's own acceptance quantity
    max|f/ewt| <= 1 (src/sparse.f:126-129) with f from the oracle's host model,
(3) a manufactured-solution run at full size (closed-form discrete steady state)."""
import os

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

C5 = int(os.environ.get("RS_TEST_C5_SIZE", "256"))


def test_config_c5_full_size(built, oracle):
    """BASELINE config C5, the workload bench.py times: ian/jan/igp/jgp at 256^3 bit-exact, ngp = 13, the returned
    state passes the reference's residual test evaluated on the host."""
    import rootsolve_b200 as rs
    S = C5
    n = S ** 3
    ipar = [1, S, S, S, 0]
    out = rs.steady_3D(np.full(n, 10.0), func="diff3d", parms=H.P_DIFF3D, dimens=[S, S, S], positive=True, ipar=ipar,
                       return_groups=True, **H.TOL)
    A = out.attributes
    assert A["steady"]
    rian, rjan = oracle.sparse_structure("3D", 1, (S, S, S), (0, 0, 0))
    assert A["dims"]["nnz"] == len(rjan) == 7 * n - 6 * S * S
    assert np.array_equal(A["ian"], rian) and np.array_equal(A["jan"], rjan)
    rngp, rigp, rjgp = oracle.jgroup(rian, rjan)
    assert A["dims"]["ngp"] == rngp
    if S == 256:
        assert rngp == 13
    assert np.array_equal(A["igp"], rigp) and np.array_equal(A["jgp"], rjgp)
    f = H.host_residual(oracle, "diff3d", H.P_DIFF3D, ipar, out["y"])
    assert H.accept(out["y"], f) <= 1.0
    # mirror symmetries of the cube, positivity, values between 0 and the imposed boundary value
    Y = out["y"].reshape(S, S, S)
    assert np.allclose(Y, Y[::-1], rtol=0, atol=1e-9) and np.allclose(Y, Y.transpose(2, 1, 0), rtol=0, atol=1e-9)
    assert out["y"].min() >= 0 and out["y"].max() <= 1.0 + 1e-9


def test_manufactured_solution_full_size(built, oracle):
    """diff3dms at the C5 grid: the discrete steady state is known in closed form; |y - y*| <= atol + rtol |y*|"""
    import rootsolve_b200 as rs
    S = C5
    d = (S, S, S)
    parms, rpar, ystar = H.ms_setup(d)
    ipar = [1, S, S, S, 0]
    out = rs.steady_3D(np.full(S ** 3, 10.0), func="diff3dms", parms=parms, dimens=[S, S, S], positive=True, ipar=ipar,
                       rpar=rpar, **H.TOL)
    assert out.attributes["steady"]
    ok, worst = H.within_tol(out["y"], ystar)
    assert ok, f"max err / (atol + rtol|y*|) = {worst}"
    f = H.host_residual(oracle, "diff3dms", parms, ipar, out["y"], rpar=rpar)
    assert H.accept(out["y"], f) <= 1.0


def test_manufactured_solution_small_matches_oracle(built, oracle):
    """the same model at a size the oracle solves: oracle, GPU and closed form agree"""
    import rootsolve_b200 as rs
    d = (12, 10, 14)
    parms, rpar, ystar = H.ms_setup(d)
    n = int(np.prod(d))
    ipar = [1, *d, 0]
    ref = oracle.stsparse(np.full(n, 10.0), oracle.model("diff3dms", parms), sparsetype="3D", nnz=[1, *d, 0, 0, 0], ipar=ipar,
                          rpar=rpar, lrw=2000000, positive=True, **H.TOL)
    out = rs.stodes(np.full(n, 10.0), func="diff3dms", parms=parms, sparsetype="3D", nnz=[1, *d, 0, 0, 0], ipar=ipar,
                    rpar=rpar, positive=True, **H.TOL)
    assert ref["steady"] and out.attributes["steady"]
    assert H.within_tol(ref["y"], ystar)[0]
    ok, worst = H.within_tol(out["y"], ref["y"])
    assert ok, worst


BIG = [
    # (model, parms, sparsetype, stodes dims, nspec, y0 per species, variant, oracle lrw)
    ("diff3d", H.P_DIFF3D, "3D", (24, 24, 24), 1, [10.0], 0, 40_000_000),
    ("diff3d", H.P_DIFF3D, "3D", (32, 32, 32), 1, [10.0], 0, 150_000_000),
    ("rd3d", H.P_RD3D, "3D", (16, 16, 16), 2, [10.0, 1.0], 0, 60_000_000),
    ("diff2d", H.P_DIFF2D_MONOD, "2D", (128, 128), 1, [10.0], 1, 8_000_000),
    ("diff2d", H.P_DIFF2D_MONOD, "2D", (256, 256), 1, [10.0], 1, 40_000_000),
]


@pytest.mark.parametrize("model,parms,st,d,nspec,y0s,variant,lrw", BIG, ids=lambda v: str(v) if isinstance(v, (str, tuple)) else None)
def test_steady_state_matches_oracle_large(built, oracle, model, parms, st, d, nspec, y0s, variant, lrw):
    """the regime where the linear tolerance and the Krylov cycle logic matter (hundreds of unknowns per direction of
    coupling), compared with the direct solve: |y_gpu - y_oracle| <= atol + rtol |y|"""
    import rootsolve_b200 as rs
    nc = int(np.prod(d))
    y0 = np.concatenate([np.full(nc, v) for v in y0s])
    pad = [1] * (3 - len(d))
    ipar = [nspec, *d, *pad, variant]
    nnz = [nspec, *d] + [0] * len(d)
    ref = oracle.stsparse(y0, oracle.model(model, parms), sparsetype=st, nnz=nnz, ipar=ipar, lrw=lrw, positive=True, **H.TOL)
    out = rs.stodes(y0, func=model, parms=parms, sparsetype=st, nnz=nnz, ipar=ipar, positive=True, return_groups=True, **H.TOL)
    A = out.attributes
    assert ref["steady"] and A["steady"]
    assert A["dims"]["nnz"] == ref["dims"]["nnz"] and A["dims"]["ngp"] == ref["dims"]["ngp"]
    assert np.array_equal(A["igp"], ref["igp"]) and np.array_equal(A["jgp"], ref["jgp"])
    ok, worst = H.within_tol(out["y"], ref["y"])
    assert ok, f"max err / (atol+rtol|y|) = {worst}"
