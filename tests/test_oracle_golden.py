"""Pins the CPU oracle (oracle/stodes_oracle.c) to everything the reference holds for this path
(SURVEY.md 8(c)): the frozen vignette output in inst/doc/rootSolve.pdf p.16, the closed-form
nonzero counts implied by src/sparsecoeff.f, the group counts of JGROUP, and an independent
sparse direct solve for the Yale LDU chain.  CPU only."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

import helpers as H

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_pdf_p16_golden(oracle):
    """nnz=6400, ngp=12, work length 82366, precis = 1.105000e+01 3.917674e-07 2.751792e-16"""
    g = json.load(open(os.path.join(GOLD, "rootsolve_pdf_p16.json")))
    n = g["n"]
    r = oracle.stsparse(np.full(n ** 3, g["y0"]), oracle.model("diff3d", g["parms"]), sparsetype="3D",
                        nnz=[1, n, n, n, 0, 0, 0], atol=g["tol"], rtol=g["tol"], ctol=g["tol"], lrw=g["lrw_given"],
                        positive=True, ipar=[1, n, n, n, 1])
    assert r["steady"] is True
    assert r["dims"] == {"nnz": g["nnz"], "ngp": g["ngp"], "lrw": g["lrw_used"]}
    assert len(r["precis"]) == 3
    # the PDF prints 7 significant digits
    for got, want in zip(r["precis"], g["precis"]):
        assert float("%.6e" % got) == want
    # R-side estimates printed by verbose=TRUE (R/stodes.R:139,143,182-183)
    Type, nnzv, ngp, lrw, itol = oracle.stodes_args(n ** 3, "3D", [1, n, n, n, 0, 0, 0])
    assert (Type, int(nnzv[0]), ngp, itol) == (4, g["nnz_estimated_by_R"], g["function_calls_estimated_by_R"], 1)


def test_pdf_p8_anoxic_model_golden(oracle):
    """The reference's compiled-code example (inst/dynload/anoxmodc.c) solved by its dense sibling stode: frozen y,
    SumS and the eight precis values (pdf p.8-9).  The oracle's Newton skeleton reproduces all of them once it always
    solves J x = beta (as stode does) and the igp terminator is repaired; run exactly as stodes would run it, the
    reference's unterminated igp (every column its own group) leaves the last Jacobian column empty -> zero pivot."""
    g = json.load(open(os.path.join(GOLD, "rootsolve_pdf_p8_anox.json")))
    f = oracle.model("anoxmod", g["parms"])
    r = oracle.stsparse(np.asarray(g["y0"]), f, sparsetype="sparseint", nnz=None, positive=True, nout=1, lrw=1000,
                        quirk_first_step_transposed=False, quirk_unterminated_igp=False)
    assert r["steady"] and r["dims"]["nnz"] == 14 and r["dims"]["ngp"] == 4
    assert len(r["precis"]) == len(g["precis"])
    # printed with 7 significant digits; the golden comes from the dense LINPACK sibling, so round-off differences
    # grow as the quadratically converging iteration approaches 1e-8 (last two values: 1e-5 and 3e-3 relative)
    for got, want, tol in zip(r["precis"], g["precis"], [2e-6] * 6 + [1e-4, 1e-2]):
        assert abs(got - want) <= tol * abs(want)
    for got, want in zip(r["y"], g["y"]):
        assert float("%.10g" % got) == pytest.approx(want, rel=1e-9, abs=5e-7)
    assert r["out"][0] == pytest.approx(g["SumS"], rel=1e-12)
    with pytest.raises(RuntimeError, match="zero pivot"):
        oracle.stsparse(np.asarray(g["y0"]), f, sparsetype="sparseint", nnz=None, positive=True, nout=1, lrw=1000)


def test_pdf_p8_golden_with_flux_and_bo2_as_forcings(oracle):
    """The same model with Flux and BO2 handed over as forcing function values through the model's initforc
    (src/call_stsparse.c:242-251, src/steady_utils.c:62-77): same golden state; wrong parameter slots are ignored."""
    g = json.load(open(os.path.join(GOLD, "rootsolve_pdf_p8_anox.json")))
    parms = list(g["parms"])
    flux, bo2 = parms[1], parms[6]
    parms[1] = parms[6] = -1.0e30          # must not be used
    f = oracle.model("anoxforc", parms, forcs=[flux, bo2])
    r = oracle.stsparse(np.asarray(g["y0"]), f, sparsetype="sparseint", nnz=None, positive=True, nout=1, lrw=1000,
                        quirk_first_step_transposed=False, quirk_unterminated_igp=False)
    assert r["steady"] and len(r["precis"]) == len(g["precis"])
    for got, want in zip(r["y"], g["y"]):
        assert float("%.10g" % got) == pytest.approx(want, rel=1e-9, abs=5e-7)
    with pytest.raises(ValueError, match="Confusion over the length of forcs"):
        oracle.model("anoxforc", parms, forcs=[flux])


def test_structure_hash_fixture(oracle):
    for e in json.load(open(os.path.join(GOLD, "structure_hashes.json"))):
        ian, jan = oracle.sparse_structure(e["kind"], e["nspec"], e["dims"], e["cyclic"])
        ngp, igp, jgp = oracle.jgroup(ian, jan)
        sha = lambda a: hashlib.sha256(np.ascontiguousarray(a, dtype=np.int32).tobytes()).hexdigest()
        assert (len(jan), ngp) == (e["nnz"], e["ngp"])
        assert (sha(ian), sha(jan), sha(igp), sha(jgp)) == (e["ian"], e["jan"], e["igp"], e["jgp"])


@pytest.mark.parametrize("nspec,nb", [(1, 10), (3, 40), (5, 1000)])
def test_nnz_closed_form_1d(oracle, nspec, nb):
    n = nspec * nb
    ian, jan = oracle.sparse_structure("1D", nspec, [nb], [0])
    assert len(jan) == n * (2 + nspec) - 2 * nspec                     # R/stodes.R:104
    ianc, janc = oracle.sparse_structure("1D", nspec, [nb], [1])
    assert len(janc) == n * (2 + nspec)                               # + 2*nspec when cyclic, R/stodes.R:106-108
    ngp, igp, jgp = oracle.jgroup(ian, jan)
    assert ngp == 3 * nspec                                            # SURVEY 8(c) closed form
    assert sorted(jgp.tolist()) == list(range(1, n + 1))


@pytest.mark.parametrize("nspec,d", [(1, (50, 50)), (1, (100, 100)), (2, (17, 9)), (1, (50, 100))])
def test_nnz_closed_form_2d(oracle, nspec, d):
    n = nspec * d[0] * d[1]
    ian, jan = oracle.sparse_structure("2D", nspec, d, [0, 0])
    assert len(jan) == n * (4 + nspec) - 2 * nspec * (d[0] + d[1])     # R/stodes.R:116-117
    ngp, _, _ = oracle.jgroup(ian, jan)
    if nspec == 1:
        assert ngp == 7


@pytest.mark.parametrize("nspec,d", [(1, (10, 10, 10)), (2, (6, 7, 8)), (1, (16, 16, 16))])
def test_nnz_closed_form_3d(oracle, nspec, d):
    n = nspec * d[0] * d[1] * d[2]
    ian, jan = oracle.sparse_structure("3D", nspec, d, [0, 0, 0])
    assert len(jan) == n * (6 + nspec) - 2 * nspec * (d[1] * d[2] + d[0] * d[2] + d[0] * d[1])
    # every column's first entry is its own diagonal (src/sparsecoeff.f:247)
    assert np.array_equal(jan[ian[:-1] - 1], np.arange(1, n + 1))


def test_grouping_is_a_valid_partition(oracle):
    """columns of one group never share a row (what makes the grouped FD Jacobian exact)"""
    ian, jan = oracle.sparse_structure("3D", 2, (5, 6, 7), (1, 0, 1))
    ngp, igp, jgp = oracle.jgroup(ian, jan)
    n = len(ian) - 1
    for g in range(ngp):
        rows = np.concatenate([jan[ian[j - 1] - 1: ian[j] - 1] for j in jgp[igp[g] - 1: igp[g + 1] - 1]])
        assert len(rows) == len(set(rows.tolist()))
    assert igp[0] == 1 and igp[-1] == n + 1


def test_perturb_rule(oracle):
    L = oracle.lib()
    for v in [0.0, 1e-12, -3.0, 10.0, 1e9]:
        x = C.c_double(v)
        d = L.orc_perturb(C.byref(x))
        assert d == max(abs(v) * 1e-8, 1e-8) and x.value == v + d


def test_yale_solve_against_scipy(oracle):
    """odrv + cdrv(path 1) solve M x = b and cdrv(path 4) solves M' x = b (src/sparse.f:136-149)"""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla
    rng = np.random.default_rng(20260101)
    L = oracle.lib()
    for kind, nspec, dims in [("2D", 1, (15, 12)), ("3D", 2, (5, 6, 4)), ("1D", 4, (60,))]:
        ian, jan = oracle.sparse_structure(kind, nspec, dims, [0] * len(dims))
        n = len(ian) - 1
        a = rng.uniform(-1, 1, len(jan))
        a[ian[:-1] - 1] = 5.0 + rng.uniform(0, 1, n)
        b = rng.standard_normal(n)
        M = sp.csr_matrix((a.copy(), jan - 1, ian - 1), shape=(n, n))   # (ian,jan,a) read by rows is M
        nsp = 40 * len(jan)
        isp = np.zeros(2 * nsp, dtype=np.int32)
        rsp = np.zeros(nsp)
        r = np.zeros(n, dtype=np.int32)
        ic = np.zeros(n, dtype=np.int32)
        ianw, janw, aw = ian.copy(), jan.copy(), a.copy()
        assert L.orc_odrv(n, H.ip(ianw), H.ip(janw), H.ip(r), H.ip(ic), nsp, H.ip(isp), 1) == 0
        assert sorted(r.tolist()) == list(range(1, n + 1))
        c = np.zeros(n, dtype=np.int32)
        c[ic - 1] = np.arange(1, n + 1)
        x = np.zeros(n)
        esp = C.c_int(0)
        assert L.orc_cdrv(n, H.ip(r), H.ip(c), H.ip(ic), H.ip(ianw), H.ip(janw), H.dp(aw), H.dp(b), H.dp(x), nsp,
                          H.ip(isp), H.dp(rsp), C.byref(esp), 1) == 0
        assert np.allclose(x, spla.spsolve(M.tocsc(), b), rtol=1e-10, atol=1e-12)
        xt = np.zeros(n)
        assert L.orc_cdrv(n, H.ip(r), H.ip(c), H.ip(ic), H.ip(ianw), H.ip(janw), H.dp(aw), H.dp(b), H.dp(xt), nsp,
                          H.ip(isp), H.dp(rsp), C.byref(esp), 4) == 0
        assert np.allclose(xt, spla.spsolve(M.T.tocsc(), b), rtol=1e-10, atol=1e-12)
        assert esp.value > 0


def test_oracle_newton_on_nonlinear_cases(oracle):
    """C1 (2-D Monod) and the two-species 3-D case converge, stay positive, and satisfy the residual test"""
    d = (50, 50)
    r = oracle.stsparse(np.full(d[0] * d[1], 10.0), oracle.model("diff2d", H.P_DIFF2D_MONOD), sparsetype="2D",
                        nnz=[1, d[0], d[1], 0, 0], ipar=[1, d[0], d[1], 1, 1], lrw=600000, positive=True, **H.TOL)
    assert r["steady"] and r["dims"]["nnz"] == 12300 and r["dims"]["ngp"] == 7
    assert r["y"].min() > 0 and r["y"].max() < 1.0
    # the transposed-first-step quirk does not move the root
    r2 = oracle.stsparse(np.full(d[0] * d[1], 10.0), oracle.model("diff2d", H.P_DIFF2D_MONOD), sparsetype="2D",
                         nnz=[1, d[0], d[1], 0, 0], ipar=[1, d[0], d[1], 1, 1], lrw=600000, positive=True,
                         quirk_first_step_transposed=False, **H.TOL)
    ok, worst = H.within_tol(r2["y"], r["y"])
    assert ok, worst
    sd = (6, 7, 8)
    nc = int(np.prod(sd))
    y0 = np.concatenate([np.full(nc, 10.0), np.full(nc, 1.0)])
    q = oracle.stsparse(y0, oracle.model("rd3d", H.P_RD3D), sparsetype="3D", nnz=[2, *sd, 0, 0, 0], ipar=[2, *sd, 0],
                        lrw=2000000, positive=True, **H.TOL)
    q2 = oracle.stsparse(y0, oracle.model("rd3d", H.P_RD3D), sparsetype="3D", nnz=[2, *sd, 0, 0, 0], ipar=[2, *sd, 0],
                         lrw=2000000, positive=True, quirk_first_step_transposed=False, **H.TOL)
    assert q["steady"] and q2["steady"]
    ok, worst = H.within_tol(q2["y"], q["y"], factor=2.0)
    assert ok, worst


def test_man_page_examples_fit_their_work_arrays(oracle):
    """The two steady.2D examples of the reference state the Yale work length they run with: lrw = 1e6 for the
    100 x 100 model (man/steady.2D.Rd:139-143) and lrw = 283800 for the 50 x 100 model with a cyclic boundary in y
    (:202-206).  The restated ordering + symbolic LDU must fit both (an upper-bound pin of md/nsfc fill), and the
    cyclic pattern gets its wrap entries and one extra group per cyclic direction (R/stodes.R:123-131)."""
    g = json.load(open(os.path.join(GOLD, "rootsolve_man_steady2d.json")))
    assert g["example1"]["lrw_given"] == 1000000 and g["example2"]["lrw_given"] == g["example2"]["lrw_needed"] == 283800
    n = g["example1"]["dimens"][0]
    a = oracle.stsparse(np.full(n * n, 10.0), oracle.model("diff2d", H.P_DIFF2D_LIN), sparsetype="2D", nnz=[1, n, n, 0, 0],
                        ipar=[1, n, n, 1, 0], lrw=g["example1"]["lrw_given"], positive=True, **H.TOL)
    assert a["steady"] and a["dims"]["nnz"] == 5 * n * n - 4 * n and a["dims"]["ngp"] == 7
    assert 0 < a["dims"]["lrw"] <= 1000000
    nx, ny = 50, 100                      # R: matrix(nrow = nx, ncol = ny); stodes sees rev(dimens) = (ny, nx), rev(Bnd) = (1, 0)
    b = oracle.stsparse(np.full(nx * ny, 10.0), oracle.model("diff2db", H.P_DIFF2DB), sparsetype="2D", nnz=[1, ny, nx, 1, 0],
                        ipar=[1, ny, nx, 1, 0], lrw=283800, positive=True, **H.TOL)
    assert b["steady"]
    assert b["dims"]["nnz"] == 5 * nx * ny - 2 * (nx + ny) + 2 * nx       # two wrap entries per line across the cyclic side
    # 283800 is not just enough, it is EXACTLY what the restated minimum-degree ordering + symbolic LDU need -- the
    # man page's author evidently copied attributes(ST2)$dims back into the call.  A second exact pin of md / nsfc /
    # the cdrv storage carving (after the 82366 of pdf p.16), on a cyclic pattern with a nonsymmetric Jacobian.
    assert b["dims"]["lrw"] == 283800
    with pytest.raises(RuntimeError, match="insufficient storage"):
        oracle.stsparse(np.full(nx * ny, 10.0), oracle.model("diff2db", H.P_DIFF2DB), sparsetype="2D", nnz=[1, ny, nx, 1, 0],
                        ipar=[1, ny, nx, 1, 0], lrw=283799, positive=True, **H.TOL)
    y = b["y"].reshape(ny, nx)
    assert y.min() > 0 and y[0].mean() > y[ny // 2].mean()                # the extra flux enters along the first column of the R matrix


def test_man_page_river_example(oracle):
    """man/steady.1D.Rd:113-153 (O2BOD), run through the stodes path: 2 species x 100 boxes, pos = TRUE, one output
    vector.  No numbers are printed in the man page; what it shows is the physics: BOD decays downstream, the
    oxygen sag recovers towards saturation, and the output vector is the decay rate at the returned state."""
    N = 100
    y0 = np.concatenate([np.full(N, 200.0), np.full(N, 200.0)])
    r = oracle.stsparse(y0, oracle.model("o2bod", H.P_O2BOD), sparsetype="1D", nnz=[2, N, 0], ipar=[2, N], lrw=100000,
                        positive=True, nout=N)
    assert r["steady"] and r["dims"]["nnz"] == 2 * N * 4 - 4 and r["dims"]["ngp"] == 6
    BOD, O2 = r["y"][:N], r["y"][N:]
    assert np.all(np.diff(BOD) < 0) and BOD[0] < 1500 and BOD[-1] > 0
    assert O2.min() < 50 and O2[-1] > O2.min() and np.all(O2 > 0)
    assert np.allclose(r["out"], 0.1 * BOD * O2 / (O2 + 10.0), rtol=1e-12)


def test_oracle_error_paths(oracle):
    with pytest.raises(RuntimeError, match="lrw"):
        oracle.stsparse(np.full(1000, 10.0), oracle.model("diff3d", H.P_DIFF3D), sparsetype="3D",
                        nnz=[1, 10, 10, 10, 0, 0, 0], ipar=[1, 10, 10, 10, 0], lrw=20000, **H.TOL)
    with pytest.raises(RuntimeError, match="not compatible"):
        oracle.stsparse(np.full(1001, 10.0), oracle.model("diff3d", H.P_DIFF3D), sparsetype="3D",
                        nnz=[2, 10, 10, 10, 0, 0, 0], ipar=[1, 10, 10, 10, 0], lrw=200000, **H.TOL)
