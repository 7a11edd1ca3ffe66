"""Host-side mirror of rootSolve's R front end for the stodes path.

Same function names, argument names, defaults and error messages as
``R/stodes.R:7-398``, ``R/steady.1D.R``, ``R/steady.2D.R``, ``R/steady.3D.R`` and
``R/steady.R`` -- for the ``method="stodes"`` branch only; everything numeric
happens behind the C ABI (``include/rootsolve_b200.h``) on the GPU.

Differences that follow from BASELINE.json's north star:
  * ``func`` is a compiled *device* model: the name of a model exported by a
    model library (``dllname``; default: the bundled ``rs_models``).  R closures
    cannot run on the GPU and are rejected.
  * every ``spmethod`` (``"yale"``, ``"ilut"``, ``"ilutp"``) is served by the preconditioned Krylov solver;
    ``lrw`` is accepted and ignored; ``control`` carries the Krylov knobs (``precond``, ``max_krylov``,
    ``lin_rtol``, ``lin_ewt_factor``, ``check_every``) and may carry the reference's ILUT options, which are ignored.
"""
from __future__ import annotations

import ctypes as C
import warnings

import numpy as np

from . import _capi

_PRECOND = {"none": _capi.RS_PRECOND_NONE, "jacobi": _capi.RS_PRECOND_JACOBI,
            "block-jacobi": _capi.RS_PRECOND_BLOCK_JACOBI, "mg": _capi.RS_PRECOND_MG, "btd": _capi.RS_PRECOND_BTD,
            "auto": _capi.RS_PRECOND_AUTO}


class SteadyResult(dict):
    """A list-with-attributes: ``res["y"]`` plus ``res.attributes`` (precis, steady, dims, ...)."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.attributes = {}
        self.rclass = ["list"]

    def attr(self, name):
        return self.attributes.get(name)


def _ip(a):
    return a.ctypes.data_as(_capi.c_int_p)


def _dp(a):
    return a.ctypes.data_as(_capi.c_double_p)


def sparse_args(n, sparsetype, nnz=None, inz=None, ngp=None, lrw=None):
    """``Type``, the ``nnz`` vector, ``ngp`` and ``lrw`` exactly as R/stodes.R:66-160 derives them."""
    Type, ian, jan = 1, None, None
    if ngp is None:
        ngp = n + 1
    if sparsetype in ("sparseint", "sparsereturn", "sparse"):
        # Type 1: sparsity estimated numerically by perturbation (R/stodes.R:71-78); nnz defaults to n*n
        nz = n * n if (nnz is None or sparsetype == "sparse") else int(np.atleast_1d(nnz)[0])
        nnzv = [min(nz, 2 ** 31 - 1)]
    elif sparsetype == "sparseusr":
        if inz is None:
            raise ValueError("'inz' must be specified if 'sparsetype' = 'sparseusr'")
        inz = np.asarray(inz)
        Type = 0
        rr = inz[:, 1]
        if len(rr) > 1 and np.min(np.diff(rr)) < 0:
            raise ValueError("cannot proceed: row indices in inz should be sorted")
        nz = inz.shape[0]
        ian = np.ones(n + 1, dtype=np.int32)
        jan = np.zeros(nz, dtype=np.int32)
        for i in range(1, n + 1):
            ii = np.nonzero(rr == i)[0]
            ian[i] = ian[i - 1] + len(ii)
            jan[ian[i - 1] - 1: ian[i] - 1] = inz[ii, 0]
        nnzv = [nz]
    elif sparsetype == "sparsejan":
        if inz is None:
            raise ValueError("'inz' must be specified if 'sparsetype' = 'sparsejan'")
        inz = np.asarray(inz, dtype=np.int32).ravel()
        Type = 0
        ian = inz[: n + 1].copy()
        jan = inz[n + 1:].copy()
        nnzv = [len(inz) - n - 1]
    elif sparsetype == "1D":
        Type = 2
        nspec = int(nnz[0])
        nnzv = [n * (2 + nspec) - 2 * nspec] + [int(v) for v in nnz]
        ngp = 3 * nspec + 1
        if nnzv[3] == 1:
            nnzv[0] += 2 * nspec
            ngp += 1
    elif sparsetype in ("2D", "2Dmap"):
        # "2Dmap": nnz = (nspec, dimens, cyclic, TotN, map[TotN]) and y holds only the states present in the map
        Type = 3 if sparsetype == "2D" else 30
        nspec = int(nnz[0])
        dimens = [int(v) for v in nnz[1:3]]
        nstate = n if Type == 3 else nspec * dimens[0] * dimens[1]     # R/stodes.R:116-119
        nnzv = [nstate * (4 + nspec) - 2 * nspec * sum(dimens)] + [int(v) for v in nnz]
        # R/stodes.R:121 "ngp < 4*nspec+1" compares instead of assigning: ngp stays n+1
        dimmax = max(dimens)
        for slot in (4, 5):
            if nnzv[slot] == 1:
                nnzv[0] += 2 * dimmax * nspec
                ngp += 1
    elif sparsetype in ("3D", "3Dmap"):
        Type = 4 if sparsetype == "3D" else 40
        nspec = int(nnz[0])
        dimens = [int(v) for v in nnz[1:4]]
        nstate = n if Type == 4 else nspec * dimens[0] * dimens[1] * dimens[2]   # R/stodes.R:138-141
        nnzv = [nstate * (6 + nspec) - 3 * nspec * sum(dimens)] + [int(v) for v in nnz]
        if nnzv[5] == 1:
            nnzv[0] += 2 * dimens[1] * dimens[2] * nspec
            ngp += 1
        if nnzv[6] == 1:
            nnzv[0] += 2 * dimens[0] * dimens[2] * nspec
            ngp += 1
        if nnzv[7] == 1:
            nnzv[0] += 2 * dimens[0] * dimens[1] * nspec
            ngp += 1
    else:
        raise ValueError("cannot run stodes: sparsetype not known ")
    if lrw is None:
        lrw = 3 * n + 4 * nnzv[0]
    return Type, np.asarray(nnzv, dtype=np.int32), int(ngp), int(min(lrw, 2**31 - 1)), ian, jan


def approx(xy, xout, method="linear", rule=1, f=0.0, ties="mean"):
    """R's ``approx(x, y, xout, method, rule, f, ties)`` at ONE point, which is all ``stodes`` asks of it
    (R/stodes.R:229-236).  rule 1: NaN outside the data range, 2: the value at the closest extreme."""
    xy = np.asarray(xy, dtype=np.float64)
    if xy.ndim != 2 or xy.shape[1] != 2:
        raise ValueError("a forcing function data set must be a two-columned matrix (time, value)")
    ok = ~(np.isnan(xy[:, 0]) | np.isnan(xy[:, 1]))
    x, v = xy[ok, 0], xy[ok, 1]
    order = np.argsort(x, kind="stable")
    x, v = x[order], v[order]
    ux, start = np.unique(x, return_index=True)
    if len(ux) < len(x):                       # tied times
        if ties not in ("mean", "ordered", "min", "max"):
            raise ValueError("ties must be 'mean', 'ordered', 'min' or 'max'")
        groups = np.split(v, start[1:])
        pick = {"mean": np.mean, "min": np.min, "max": np.max, "ordered": lambda g: g[-1]}[ties]
        x, v = ux, np.asarray([pick(g) for g in groups])
    if len(x) == 0 or (method == "linear" and len(x) < 2):
        raise ValueError("need at least two non-NA values to interpolate")
    rules = (rule, rule) if np.ndim(rule) == 0 else tuple(rule)
    if xout < x[0]:
        return float(v[0]) if rules[0] == 2 else float("nan")
    if xout > x[-1]:
        return float(v[-1]) if rules[1] == 2 else float("nan")
    hi = int(np.searchsorted(x, xout, side="left"))
    if x[hi] == xout:
        return float(v[hi])
    lo = hi - 1
    if method == "linear":
        return float(v[lo] + (v[hi] - v[lo]) * ((xout - x[lo]) / (x[hi] - x[lo])))
    if method == "constant":
        return float(v[lo] * (1.0 - f) + v[hi] * f) if f != 0.0 else float(v[lo])
    raise ValueError("approx: invalid interpolation method")


def forcing_values(forcings, time, fcontrol=None):
    """The ``Forc`` vector of R/stodes.R:229-236: a list of (time, value) data sets is interpolated at `time`;
    anything else is taken as the vector of forcing function values itself."""
    if isinstance(forcings, (list, tuple)) and len(forcings) > 0 and all(np.ndim(m) == 2 for m in forcings):
        return np.asarray([approx(m, time, **(fcontrol or {})) for m in forcings], dtype=np.float64)
    return np.ascontiguousarray(np.atleast_1d(np.asarray(forcings, dtype=np.float64)).ravel())


def _itol(n, atol, rtol):
    itol = 1
    if len(atol) == n and len(rtol) == n:
        itol = 4
    elif len(atol) == n and len(rtol) != n:
        itol = 2
    elif len(atol) != n and len(rtol) == n:
        itol = 3
    return itol


def _control(control):
    if control is None:
        return None
    unknown = set(control) - {"precond", "max_krylov", "lin_rtol", "lin_ewt_factor", "check_every",
                              "droptol", "permtol", "fillin", "lenplufac"}   # the last four: the reference's ILUT knobs
    if unknown:
        raise ValueError("unknown control option(s): %s" % sorted(unknown))
    kc = _capi.KrylovControl(precond=_capi.RS_PRECOND_AUTO)
    pc = control.get("precond", "auto")
    if pc not in _PRECOND:
        raise ValueError("control$precond must be one of %s" % sorted(_PRECOND))
    kc.precond = _PRECOND[pc]
    kc.max_krylov = int(control.get("max_krylov", 0))
    kc.lin_rtol = float(control.get("lin_rtol", 0.0))
    kc.lin_ewt_factor = float(control.get("lin_ewt_factor", 0.0))
    kc.check_every = int(control.get("check_every", 0))
    return kc


class _Packed:
    """Keeps the numpy buffers behind an rs_stsparse_args alive."""

    def __init__(self):
        self.keep = []
        self.args = _capi.StsparseArgs()


def pack_args(y, time, func, parms, rtol, atol, ctol, sparsetype, verbose, nnz, inz, lrw, ngp, positive, maxiter,
              dllname, initfunc, initpar, rpar, ipar, nout, spmethod, control):
    y = np.array(y, dtype=np.float64)
    if y.dtype.kind not in "fiu":
        raise TypeError("`y' must be numeric")
    y = np.ascontiguousarray(y.ravel(order="F"))
    n = len(y)
    if not np.isscalar(maxiter) or int(maxiter) < 1:
        raise ValueError("maxiter must be >=1")
    atol = np.atleast_1d(np.asarray(atol, dtype=np.float64))
    rtol = np.atleast_1d(np.asarray(rtol, dtype=np.float64))
    if len(atol) > 1 and len(atol) != n:
        raise ValueError("`atol' must either be a scalar, or as long as `y'")
    if len(rtol) > 1 and len(rtol) != n:
        raise ValueError("`rtol' must either be a scalar, or as long as `y'")
    if np.ndim(ctol) > 0 and np.size(ctol) > 1:
        raise ValueError("`ctol' must be a scalar")
    if callable(func):
        raise TypeError("`func' must be the name of a compiled device model (see rootsolve_b200_model.cuh); "
                        "R/Python closures cannot run on the GPU path")
    if not isinstance(func, (str, int)):
        raise TypeError("`func' must be a function or character vector or a compiled function")
    # R/stodes.R:314-329.  In the reference "ilut"/"ilutp" swap the Yale LDU for SPARSKIT's incomplete LU, applied
    # once per Newton step as if it were exact (src/sparsekit.f:141-166).  Here every spmethod is served by the same
    # preconditioned Krylov solve: same tests, same steady state, fewer and more accurate Newton steps.  The ILUT
    # knobs of `control' (droptol, permtol, fillin, lenplufac; R/stodes.R:407-420) are accepted; nothing uses them.
    if spmethod not in ("yale", "ilut", "ilutp"):
        raise ValueError("`spmethod' must be one of `yale', `ilut' or `ilutp'")
    ret_only = sparsetype == "sparsereturn"
    Type, nnzv, maxg, lrw, ian, jan = sparse_args(n, sparsetype, nnz, inz, ngp, lrw)
    imethod = {"yale": 1, "ilut": 2, "ilutp": 3}[spmethod]
    if isinstance(positive, (bool, np.bool_)):
        Pos, positive_v = int(positive), np.asarray([int(positive)], dtype=np.int32)
    else:
        positive_v = np.atleast_1d(np.asarray(positive, dtype=np.int32))
        if positive_v.max() > n:
            raise ValueError("the elements of 'positive' should be < the number of state variables")
        if positive_v.min() < 1:
            raise ValueError("the elements of 'positive' should be >0")
        Pos = 0
    if initfunc is None:
        initpar = None
    sym = func if isinstance(func, int) else _capi.model_symbol(func, dllname)
    P = _Packed()
    a = P.args
    initpar_a = None if initpar is None else np.ascontiguousarray(initpar, dtype=np.float64)
    rpar_a = np.zeros(0) if rpar is None else np.ascontiguousarray(rpar, dtype=np.float64)
    ipar_a = np.zeros(0, dtype=np.int32) if ipar is None else np.ascontiguousarray(ipar, dtype=np.int32)
    kc = _control(control)
    P.keep += [y, atol, rtol, nnzv, positive_v, initpar_a, rpar_a, ipar_a, ian, jan, kc]
    a.y, a.neq, a.time = _dp(y), n, float(time)
    a.func = sym
    if initpar_a is not None:
        a.initpar, a.ninitpar = _dp(initpar_a), len(initpar_a)
    a.ctol = float(ctol)
    a.atol, a.natol, a.rtol, a.nrtol = _dp(atol), len(atol), _dp(rtol), len(rtol)
    a.itol = _itol(n, atol, rtol)
    a.verbose = int(bool(verbose))
    a.nnz, a.nnnz = _ip(nnzv), len(nnzv)
    a.lrw, a.ngp, a.maxiter = lrw, maxg, int(maxiter)
    a.Pos, a.positive, a.npositive = Pos, _ip(positive_v), len(positive_v)
    a.nout = int(nout)
    a.rpar, a.nrpar = _dp(rpar_a), len(rpar_a)
    a.ipar, a.nipar = _ip(ipar_a), len(ipar_a)
    a.Type = -Type if ret_only else Type
    if ian is not None:
        a.ian, a.jan = _ip(ian), _ip(jan)
    a.imethod = 10 if ret_only else imethod
    if kc is not None:
        a.control = C.pointer(kc)
    P.y, P.n, P.maxg, P.nnz_cap = y, n, maxg, int(nnzv[0])
    return P


def stodes(y, time=0, func=None, parms=None, rtol=1e-6, atol=1e-8, ctol=1e-8, sparsetype="sparseint",
           verbose=False, nnz=None, inz=None, lrw=None, ngp=None, positive=False, maxiter=100, ynames=True,
           dllname=None, initfunc="__dllname__", initpar="__parms__", rpar=None, ipar=None, nout=0, outnames=None,
           forcings=None, initforc=None, fcontrol=None, spmethod="yale", control=None, times="__time__",
           return_groups=False):
    """Steady state of an ODE system by sparse Newton-Raphson: rootSolve's ``stodes`` (R/stodes.R:7-398)."""
    if initfunc == "__dllname__":
        initfunc = dllname if dllname is not None else "rs_models"
    if isinstance(initpar, str) and initpar == "__parms__":
        initpar = parms
    if isinstance(times, str) and times == "__time__":
        times = time
    if times is not None and np.size(times) > 1:
        warnings.warn("`times' should be one number - taking first value")
    time = 0.0 if times is None else float(np.atleast_1d(times)[0])
    P = pack_args(y, time, func, parms, rtol, atol, ctol, sparsetype, verbose, nnz, inz, lrw, ngp, positive, maxiter,
                  dllname, initfunc, initpar, rpar, ipar, nout, spmethod, control)
    if initforc is not None and forcings is not None:     # R/stodes.R:222-237: forcings without initforc are ignored
        forc = forcing_values(forcings, time, fcontrol)
        P.keep.append(forc)
        P.args.forcs, P.args.nforcs = _dp(forc), len(forc)
    n = P.n
    yout = np.zeros(n + int(nout))
    precis = np.zeros(int(maxiter))
    ian = np.zeros(n + 1, dtype=np.int32)
    ret_only = sparsetype == "sparsereturn"
    # jan is only handed back for sparsereturn / return_groups (nnz_cap is n*n for numerically detected structures)
    jan = np.zeros(P.nnz_cap if (ret_only or return_groups) else 1, dtype=np.int32)
    R = _capi.StsparseResult()
    R.y, R.precis = _dp(yout), _dp(precis)
    ret_only = sparsetype == "sparsereturn"
    if ret_only:
        R.ian, R.jan = _ip(ian), _ip(jan)
    if return_groups:
        igp = np.zeros(P.maxg + 1, dtype=np.int32)
        jgp = np.zeros(n, dtype=np.int32)
        R.igp, R.jgp = _ip(igp), _ip(jgp)
        if not ret_only:
            R.ian, R.jan = _ip(ian), _ip(jan)
    _capi.check(_capi.lib().rs_stsparse(C.byref(P.args), C.byref(R)))
    steady = [int(v) for v in R.steady]
    out = SteadyResult()
    yv = yout[:n].copy()
    out["y"] = yv
    if nout > 0:
        var = yout[n:]
        names = list(outnames) if outnames is not None else []
        names = (names + [str(i + 1) for i in range(len(names), nout)])[:nout]
        seen = []
        for nm in names:
            if nm not in seen:
                seen.append(nm)
        for nm in seen:
            out[nm] = var[[i for i, q in enumerate(names) if q == nm]].copy()
    out.attributes["precis"] = precis[: R.niter].copy()
    out.attributes["steady"] = steady[0] == 1
    if not steady[0] and not ret_only:
        warnings.warn("steady-state not reached")
    out.attributes["dims"] = {"nnz": steady[1], "ngp": steady[2], "lrw": None if steady[3] < 0 else steady[3]}
    if ret_only or return_groups:
        out.attributes["ian"] = ian
        out.attributes["jan"] = jan[: steady[1]].copy()
    if return_groups:
        out.attributes["igp"] = igp[: steady[2] + 1].copy()
        out.attributes["jgp"] = jgp
    out.attributes["krylov_iters"] = R.krylov_iters
    out.attributes["kernel_launches"] = R.kernel_launches
    out.attributes["spmv_kernel"] = {0: "csr-subwarp", 1: "csr-tma", 2: "grid-implicit-columns"}.get(R.spmv_kernel)
    out.attributes["precond"] = {0: "none", 1: "jacobi", 2: "block-jacobi", 3: "mg", 4: "btd"}.get(R.precond)
    out.attributes["krylov_method"] = {0: "bicgstab", 1: "cg"}.get(R.krylov_method)
    out.attributes["timing"] = {"setup_s": R.t_setup_s, "newton_s": R.t_newton_s}
    if verbose:
        print("precision at each steady state step")
        print(out.attributes["precis"])
        print(" nnz", steady[1], " ngp", steady[2], " device work (doubles)", steady[3])
    return out


def _bnd(cyclicBnd, nd, who):
    bnd = [0] * nd
    if cyclicBnd is not None:
        cb = np.atleast_1d(cyclicBnd)
        if cb.max() > nd:
            raise ValueError(f"cannot run {who}: cyclicBnd should be a vector or number not exceeding {nd}")
        for v in cb:
            if v > 0:
                bnd[int(v) - 1] = 1
    return bnd


def steady_1D(y, time=None, func=None, parms=None, nspec=None, dimens=None, names=None, method="stodes",
              jactype=None, cyclicBnd=None, bandwidth=1, times="__time__", **kw):
    """``steady.1D(..., method="stodes")`` (R/steady.1D.R:8-106); the banded ``stode`` branch is not this path."""
    if isinstance(times, str):
        times = time
    if dimens is None and nspec is None:
        raise ValueError("cannot run steady.1D: either nspec or dimens should be specified")
    N = int(np.size(y))
    if nspec is None:
        nspec = N // int(dimens)
    elif N % nspec != 0:
        raise ValueError("cannot run steady.1D: nspec is not an integer fraction of number of state variables")
    if names is not None and len(names) != nspec:
        raise ValueError("length of 'names' should equal 'nspec'")
    if cyclicBnd is not None:
        method = "stodes"
    if method != "stodes":
        raise NotImplementedError("only method = 'stodes' is on the GPU path (stode/runsteady are other solvers)")
    if bandwidth != 1:
        raise ValueError("cannot combine 'method = stodes' with 'bandwidth' not = 1")
    if times is None:
        times = 0
    if jactype is None:
        jactype = "1D"
    dimens = N // nspec
    Bnd = 0
    if cyclicBnd is not None:
        if np.max(cyclicBnd) > 1:
            raise ValueError("cannot run steady.1D: cyclicBnd should be NULL or a value not exceeding 1")
        Bnd = 1
    if jactype == "1D":
        out = stodes(y=y, time=times, func=func, parms=parms, nnz=[nspec, dimens, Bnd], sparsetype=jactype, **kw)
    else:
        out = stodes(y=y, time=times, func=func, parms=parms, sparsetype=jactype, **kw)
    out.rclass = ["steady1D", "rootSolve", "list"]
    if names is not None:
        out["y"] = out["y"].reshape((dimens, nspec), order="F")
    out.attributes.update(dimens=dimens, nspec=nspec, ynames=names)
    return out


def steady_2D(y, time=0, func=None, parms=None, nspec=None, dimens=None, names=None, method="stodes",
              jactype=None, cyclicBnd=None, times="__time__", **kw):
    """``steady.2D`` (R/steady.2D.R:8-66)."""
    if isinstance(times, str):
        times = time
    if dimens is None:
        raise ValueError("cannot run steady.2D: dimens should be specified")
    if len(dimens) != 2:
        raise ValueError("cannot run steady.2D: dimens should contain 2 values")
    N = int(np.size(y))
    if N % int(np.prod(dimens)) != 0:
        raise ValueError("cannot run steady.2D: dimensions are not an integer fraction of number of state variables")
    if nspec is None:
        nspec = N // int(np.prod(dimens))
    elif nspec * int(np.prod(dimens)) != N:
        raise ValueError("cannot run steady.2D: dimens[1]*dimens[2]*nspec is not equal to number of state variables")
    if names is not None and len(names) != nspec:
        raise ValueError("length of 'names' should equal 'nspec'")
    if method != "stodes":
        raise NotImplementedError("only method = 'stodes' is on the GPU path")
    if times is None:
        times = 0
    Bnd = _bnd(cyclicBnd, 2, "steady.2D")
    if jactype is None:
        jactype = "2D"
    dl = [int(v) for v in dimens]
    if jactype == "2D":  # note: stodes expects rev(dimens)
        out = stodes(y=y, time=times, func=func, parms=parms, nnz=[nspec] + dl[::-1] + Bnd[::-1], sparsetype=jactype, **kw)
    else:
        out = stodes(y=y, time=times, func=func, parms=parms, sparsetype=jactype, **kw)
    out.rclass = ["steady2D", "rootSolve", "list"]
    out.attributes.update(dimens=dl, nspec=nspec, ynames=names)
    return out


def steady_3D(y, time=0, func=None, parms=None, nspec=None, dimens=None, names=None, method="stodes",
              jactype=None, cyclicBnd=None, times="__time__", **kw):
    """``steady.3D`` (R/steady.3D.R:8-66)."""
    if isinstance(times, str):
        times = time
    if dimens is None:
        raise ValueError("cannot run steady.3D: dimens should be specified")
    if len(dimens) != 3:
        raise ValueError("cannot run steady.3D: dimens should contain 3 values")
    N = int(np.size(y))
    if N % int(np.prod(dimens)) != 0:
        raise ValueError("cannot run steady.3D: dimensions are not an integer fraction of number of state variables")
    if nspec is None:
        nspec = N // int(np.prod(dimens))
    elif nspec * int(np.prod(dimens)) != N:
        raise ValueError("cannot run steady.3D: dimens[1]*dimens[2]*dimens[3]*nspec is not equal to number of state variables")
    if names is not None and len(names) != nspec:
        raise ValueError("length of 'names' should equal 'nspec'")
    if method != "stodes":
        raise NotImplementedError("only method = 'stodes' is on the GPU path")
    if times is None:
        times = 0
    Bnd = _bnd(cyclicBnd, 3, "steady.3D")
    if jactype is None:
        jactype = "3D"
    dl = [int(v) for v in dimens]
    if jactype == "3D":  # note: stodes expects rev(dimens)
        out = stodes(y=y, times=times, func=func, parms=parms, nnz=[nspec] + dl[::-1] + Bnd[::-1], sparsetype=jactype, **kw)
    else:
        out = stodes(y=y, time=times, func=func, parms=parms, sparsetype=jactype, **kw)
    out.rclass = ["steady3D", "rootSolve", "list"]
    out.attributes.update(dimens=dl, nspec=nspec, ynames=names)
    return out


def steady(y, time=None, func=None, parms=None, method="stodes", times="__time__", **kw):
    """``steady(..., method="stodes")`` (R/steady.R:7-29)."""
    if isinstance(times, str):
        times = time
    if method != "stodes":
        raise NotImplementedError("only method = 'stodes' is on the GPU path")
    if times is None:
        times = 0
    out = stodes(y, times, func, parms=parms, **kw)
    out.rclass = ["steady", "rootSolve", "list"]
    return out
