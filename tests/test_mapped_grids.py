"""Mapped (masked) grids: sparsetype "2Dmap"/"3Dmap" (R/stodes.R:111-141, src/testMap.f:25-227,
src/call_stsparse.c:196-206).  Only the states listed in a presence map are unknowns; the Jacobian pattern is
sparse2d's / sparse3d's restricted to the present states and renumbered.

The reference ships no example or golden for these two types (nothing in man/, vignettes/ or inst/ calls them), so
the oracle is pinned here through what the code implies: with every state present the mapped generators must
reproduce sparse2d / sparse3d (which ARE pinned, test_oracle_golden.py), a mask must give exactly the sub-pattern of
the full pattern, and a solve on a full map must walk the same Newton path as the unmapped solve.  The GPU tests then
compare the CUDA path with the oracle: integer arrays bit-exact, states within atol/rtol.
"""
import ctypes as C

import numpy as np
import pytest

import helpers as H

MAP_GRIDS_2D = [(1, (1, 1), (0, 0)), (1, (7, 9), (0, 0)), (2, (17, 9), (0, 0)), (1, (12, 12), (1, 0)), (1, (12, 12), (0, 1)),
                (2, (10, 10), (1, 1)), (1, (9, 14), (0, 1)), (2, (14, 9), (0, 1)), (3, (6, 5), (0, 0))]
MAP_GRIDS_3D = [(1, (1, 1, 1), (0, 0, 0)), (1, (6, 7, 8), (0, 0, 0)), (2, (4, 5, 6), (0, 1, 0)), (1, (6, 6, 6), (1, 1, 1)),
                (1, (2, 2, 9), (1, 1, 1)), (3, (5, 5, 5), (0, 0, 0))]


def _masks(dims, rng):
    yield "full", np.ones(dims, dtype=bool)
    yield "ball", H.ball_mask(dims)
    yield "random", rng.random(dims) < 0.6
    m = np.ones(dims, dtype=bool)
    m[(0,) * len(dims)] = False          # first state absent: numbering shifts from the very start
    m[tuple(d - 1 for d in dims)] = False
    yield "corners", m


def _subpattern(ian_full, jan_full, pres):
    """The pattern of the present states, renumbered: what sparse2dmap/sparse3dmap must produce when the entry
    order of the full generator is kept."""
    ian, jan = [1], []
    for M in np.nonzero(pres > 0)[0]:
        rows = jan_full[ian_full[M] - 1: ian_full[M + 1] - 1]
        inside = (rows >= 1) & (rows <= len(pres))
        keep = np.where(inside, pres[np.clip(rows - 1, 0, len(pres) - 1)], 0)
        jan.extend(keep[keep > 0].tolist())
        ian.append(len(jan) + 1)
    return np.asarray(ian, dtype=np.int32), np.asarray(jan, dtype=np.int32)


@pytest.mark.parametrize("nspec,dims,cyc", MAP_GRIDS_2D)
def test_oracle_2dmap_is_the_subpattern_of_sparse2d(oracle, nspec, dims, cyc):
    rng = np.random.default_rng(20260101)
    # sparse2d's k==1 wrap (src/sparsecoeff.f:159 = src/testMap.f:97) leaves the species block when d1 > d2
    # (that row number is beyond the map for the last species: absent by definition here, see put_mapped)
    ian_f, jan_f = oracle.sparse_structure("2D", nspec, dims, cyc)
    for name, mask in _masks(dims, rng):
        pres, _ = H.make_map(mask, nspec)
        if pres.max() == 0:
            continue
        ian, jan = oracle.mapped_structure("2Dmap", nspec, dims, pres, cyc)
        want_ian, want_jan = _subpattern(ian_f, jan_f, pres)
        assert np.array_equal(ian, want_ian), name
        assert np.array_equal(jan, want_jan), name
        if name == "full" and jan_f.max() <= len(pres):
            assert np.array_equal(ian, ian_f) and np.array_equal(jan, jan_f)


@pytest.mark.parametrize("nspec,dims,cyc", MAP_GRIDS_3D)
def test_oracle_3dmap_is_the_subpattern_of_sparse3d(oracle, nspec, dims, cyc):
    rng = np.random.default_rng(20260102)
    ian_f, jan_f = oracle.sparse_structure("3D", nspec, dims, cyc)
    for name, mask in _masks(dims, rng):
        pres, _ = H.make_map(mask, nspec)
        if pres.max() == 0:
            continue
        ian, jan = oracle.mapped_structure("3Dmap", nspec, dims, pres, cyc)
        want_ian, want_jan = _subpattern(ian_f, jan_f, pres)
        assert np.array_equal(ian, want_ian), name
        assert np.array_equal(jan, want_jan), name
        if name == "full":
            assert np.array_equal(ian, ian_f) and np.array_equal(jan, jan_f)


def test_oracle_map_capacity_rule(oracle):
    """updatejan (src/testMap.f:14-21) stops as soon as the write position passes nnz: a capacity equal to the
    number of nonzeros is already 'not enough room'."""
    dims = (6, 7)
    pres, _ = H.make_map(H.ball_mask(dims), 1)
    ian, jan = oracle.mapped_structure("2Dmap", 1, dims, pres)
    z = len(jan)
    with pytest.raises(RuntimeError, match="rc=-2"):
        oracle.mapped_structure("2Dmap", 1, dims, pres, cap=z)
    ian2, jan2 = oracle.mapped_structure("2Dmap", 1, dims, pres, cap=z + 1)
    assert np.array_equal(jan2, jan) and np.array_equal(ian2, ian)
    with pytest.raises(RuntimeError, match="rc=-1"):   # N and nspec not compatible (src/testMap.f:51-54)
        odd = np.zeros(2 * dims[0] * dims[1], dtype=np.int32)
        odd[:3] = [1, 2, 3]
        oracle.mapped_structure("2Dmap", 2, dims, odd, cap=1000)


def _map_nnz(nspec, dims, cyc, pres):
    return [nspec] + list(dims) + list(cyc) + [len(pres)] + pres.tolist()


def test_oracle_full_map_solve_equals_unmapped_solve(oracle):
    """With every cell present, `diffmap` is `diff2d` (same arithmetic order) and sparse2dmap is sparse2d: the two
    oracle solves must agree to round-off, iteration for iteration."""
    d = (9, 11)
    n = d[0] * d[1]
    pres, inv = H.make_map(np.ones(d, dtype=bool))
    y0 = np.full(n, 10.0)
    ref = oracle.stsparse(y0, oracle.model("diff2d", H.P_DIFF2D_LIN), sparsetype="2D", nnz=[1, d[0], d[1], 0, 0],
                          ipar=[1, d[0], d[1], 1, 0], lrw=200000, positive=True, **H.TOL)
    args = dict(sparsetype="2Dmap", nnz=_map_nnz(1, d, (0, 0), pres), ipar=H.diffmap_ipar((1, d[0], d[1]), pres, inv),
                lrw=200000, positive=True, **H.TOL)
    # R sizes jan for the FULL grid (R/stodes.R:119) and updatejan wants one slot to spare: a map without a single
    # absent state is "not enough room" in the reference
    with pytest.raises(RuntimeError, match="not enough room for nonzeros"):
        oracle.stsparse(y0, oracle.model("diffmap", H.P_DIFF3D), **args)
    got = oracle.stsparse(y0, oracle.model("diffmap", H.P_DIFF3D), nnz_cap=ref["dims"]["nnz"] + 1, **args)
    assert got["steady"] and ref["steady"]
    assert np.array_equal(got["ian"], ref["ian"]) and np.array_equal(got["jan"], ref["jan"])
    assert np.array_equal(got["igp"], ref["igp"]) and np.array_equal(got["jgp"], ref["jgp"])
    assert len(got["precis"]) == len(ref["precis"])
    assert np.allclose(got["y"], ref["y"], rtol=1e-13, atol=0)


@pytest.mark.parametrize("kind,dims", [("2Dmap", (24, 31)), ("3Dmap", (9, 10, 11))])
def test_oracle_masked_solve_reaches_steady_state(oracle, kind, dims):
    """a lake / a ball: the oracle's solve on the masked grid satisfies the reference's own acceptance test"""
    pres, inv = H.make_map(H.ball_mask(dims, 0.45))
    n = len(inv)
    d3 = (1,) + tuple(dims) if kind == "2Dmap" else tuple(dims)
    ipar = H.diffmap_ipar(d3, pres, inv)
    out = oracle.stsparse(np.full(n, 10.0), oracle.model("diffmap", H.P_DIFF3D), sparsetype=kind,
                          nnz=_map_nnz(1, dims, (0,) * len(dims), pres), ipar=ipar, lrw=400000, positive=True, **H.TOL)
    assert out["steady"] and out["dims"]["nnz"] < n * (2 * len(dims) + 1)
    f = _host_residual(oracle, "diffmap", H.P_DIFF3D, ipar, out["y"])
    assert np.max(np.abs(f) / (1e-10 * np.abs(out["y"]) + 1e-10)) <= 1.0
    assert out["y"].min() > 0 and out["y"].max() < 1.0     # consumption pulls the interior below the boundary value


def _host_residual(oracle, model, parms, ipar, y):
    func = C.cast(oracle.model(model, parms), oracle.DERIV)
    n = len(y)
    f = np.zeros(n)
    ipfull = np.concatenate([[0, 0, 3 + len(ipar)], ipar]).astype(np.int32)
    neq, t = C.c_int(n), C.c_double(0.0)
    func(C.byref(neq), C.byref(t), H.dp(np.ascontiguousarray(y)), H.dp(f), H.dp(np.zeros(1)), H.ip(ipfull))
    return f


# ---- the CUDA path against the oracle ----------------------------------------------------------------------------
def gpu_map_structure(lib, kind, nspec, dims, cyc, pres, n=None, cap=None):
    from rootsolve_b200 import _capi
    t = {"2Dmap": _capi.RS_TYPE_2DMAP, "3Dmap": _capi.RS_TYPE_3DMAP}[kind]
    dims = np.asarray(dims, dtype=np.int32)
    cyc = np.asarray(cyc, dtype=np.int32)
    pres = np.ascontiguousarray(pres, dtype=np.int32)
    n = int((pres > 0).sum()) if n is None else n
    cap = len(pres) * (2 * len(dims) + nspec) + 1 if cap is None else cap
    ian = np.zeros(n + 1, dtype=np.int32)
    jan = np.zeros(cap, dtype=np.int32)
    nnz = C.c_int(cap)
    _capi.check(lib.rs_sparse_structure_map(t, n, nspec, H.ip(dims), H.ip(cyc), H.ip(pres), len(pres), C.byref(nnz), H.ip(ian),
                                            H.ip(jan)))
    return ian, jan[: nnz.value].copy()


@pytest.mark.gpu
@pytest.mark.parametrize("nspec,dims,cyc", MAP_GRIDS_2D + [(1, (300, 257), (0, 0))])
def test_gpu_2dmap_structure_bit_exact(built, oracle, nspec, dims, cyc):
    rng = np.random.default_rng(20260103)
    for name, mask in _masks(dims, rng):
        pres, _ = H.make_map(mask, nspec)
        if pres.max() == 0:
            continue
        ian_o, jan_o = oracle.mapped_structure("2Dmap", nspec, dims, pres, cyc)
        ian_g, jan_g = gpu_map_structure(built, "2Dmap", nspec, dims, cyc, pres)
        assert np.array_equal(ian_g, ian_o), name
        assert np.array_equal(jan_g, jan_o), name


@pytest.mark.gpu
@pytest.mark.parametrize("nspec,dims,cyc", MAP_GRIDS_3D + [(1, (40, 33, 37), (0, 0, 0)), (2, (12, 20, 16), (1, 0, 1))])
def test_gpu_3dmap_structure_bit_exact(built, oracle, nspec, dims, cyc):
    rng = np.random.default_rng(20260104)
    for name, mask in _masks(dims, rng):
        pres, _ = H.make_map(mask, nspec)
        if pres.max() == 0:
            continue
        ian_o, jan_o = oracle.mapped_structure("3Dmap", nspec, dims, pres, cyc)
        ian_g, jan_g = gpu_map_structure(built, "3Dmap", nspec, dims, cyc, pres)
        assert np.array_equal(ian_g, ian_o), name
        assert np.array_equal(jan_g, jan_o), name


@pytest.mark.gpu
def test_gpu_map_error_paths(built, oracle):
    from rootsolve_b200 import RootSolveError
    dims = (6, 7)
    pres, _ = H.make_map(H.ball_mask(dims), 1)
    _, jan = oracle.mapped_structure("2Dmap", 1, dims, pres)
    with pytest.raises(RootSolveError, match="not enough room for nonzeros"):      # capacity == count, as updatejan
        gpu_map_structure(built, "2Dmap", 1, dims, (0, 0), pres, cap=len(jan))
    gpu_map_structure(built, "2Dmap", 1, dims, (0, 0), pres, cap=len(jan) + 1)
    odd = np.zeros(2 * dims[0] * dims[1], dtype=np.int32)
    odd[:3] = [1, 2, 3]
    with pytest.raises(RootSolveError, match="N and nspec not compatible"):
        gpu_map_structure(built, "2Dmap", 2, dims, (0, 0), odd)
    # maps the reference would turn into garbage are refused: not in grid order, numbers repeated or skipped
    swapped = pres.copy()
    a, b = np.nonzero(pres > 0)[0][:2]
    swapped[a], swapped[b] = pres[b], pres[a]
    repeated = pres.copy()
    repeated[b] = pres[a]
    for bad, n in [(swapped, None), (repeated, None), (pres, int((pres > 0).sum()) - 1), (pres[:-1], None)]:
        with pytest.raises(RootSolveError, match="mapped grid"):
            gpu_map_structure(built, "2Dmap", 1, dims, (0, 0), bad, n=n)


@pytest.mark.gpu
@pytest.mark.parametrize("kind,dims,shrink", [("2Dmap", (24, 31), 0.45), ("2Dmap", (64, 50), 0.5), ("3Dmap", (9, 10, 11), 0.45),
                                              ("3Dmap", (20, 18, 16), 0.5)])
def test_gpu_masked_solve_matches_oracle(built, oracle, kind, dims, shrink):
    """stodes(sparsetype = "2Dmap"/"3Dmap") through the C ABI against the oracle: nnz, ngp and the four index arrays
    bit-exact, the steady state within atol/rtol."""
    import rootsolve_b200 as rs
    pres, inv = H.make_map(H.ball_mask(dims, shrink))
    n = len(inv)
    d3 = (1,) + tuple(dims) if kind == "2Dmap" else tuple(dims)
    ipar = H.diffmap_ipar(d3, pres, inv)
    nnz = _map_nnz(1, dims, (0,) * len(dims), pres)
    y0 = np.full(n, 10.0)
    ref = oracle.stsparse(y0, oracle.model("diffmap", H.P_DIFF3D), sparsetype=kind, nnz=nnz, ipar=ipar, lrw=4000000,
                          positive=True, **H.TOL)
    got = rs.stodes(y0, func="diffmap", parms=H.P_DIFF3D, sparsetype=kind, nnz=nnz, ipar=ipar, positive=True,
                    return_groups=True, **H.TOL)
    assert got.attributes["steady"] and ref["steady"]
    assert got.attributes["dims"]["nnz"] == ref["dims"]["nnz"] and got.attributes["dims"]["ngp"] == ref["dims"]["ngp"]
    # (jan itself is compared by the structure tests above: after a full solve the oracle's copy has been re-sorted
    # in place by nroc, src/sparse.f:1701-1936, as the reference's is)
    assert np.array_equal(got.attributes["ian"], ref["ian"])
    assert np.array_equal(got.attributes["jan"], oracle.mapped_structure(kind, 1, dims, pres)[1])
    assert np.array_equal(got.attributes["igp"], ref["igp"]) and np.array_equal(got.attributes["jgp"], ref["jgp"])
    assert abs(got.attributes["precis"][0] - ref["precis"][0]) <= 1e-12 * ref["precis"][0]
    ok, worst = H.within_tol(got["y"], ref["y"])
    assert ok, worst


@pytest.mark.gpu
def test_gpu_full_map_is_not_enough_room_as_in_the_reference(built):
    import rootsolve_b200 as rs
    d = (9, 11)
    pres, inv = H.make_map(np.ones(d, dtype=bool))
    with pytest.raises(rs.RootSolveError, match="not enough room for nonzeros"):
        rs.stodes(np.full(d[0] * d[1], 10.0), func="diffmap", parms=H.P_DIFF3D, sparsetype="2Dmap",
                  nnz=_map_nnz(1, d, (0, 0), pres), ipar=H.diffmap_ipar((1, d[0], d[1]), pres, inv))


@pytest.mark.gpu
def test_gpu_masked_solve_large_ball(built, oracle):
    """3Dmap at a size the direct solver cannot reach (a ball in 128^3, ~1.1 M unknowns): the reference's own
    acceptance quantity max|f/ewt| <= 1 evaluated with the HOST model on the GPU result."""
    import rootsolve_b200 as rs
    dims = (128, 128, 128)
    pres, inv = H.make_map(H.ball_mask(dims, 0.5))
    n = len(inv)
    ipar = H.diffmap_ipar(dims, pres, inv)
    out = rs.stodes(np.full(n, 10.0), func="diffmap", parms=H.P_DIFF3D, sparsetype="3Dmap",
                    nnz=_map_nnz(1, dims, (0, 0, 0), pres), ipar=ipar, positive=True, **H.TOL)
    assert out.attributes["steady"] and n > 1000000
    f = _host_residual(oracle, "diffmap", H.P_DIFF3D, ipar, out["y"])
    assert np.max(np.abs(f) / (1e-10 * np.abs(out["y"]) + 1e-10)) <= 1.0
