"""ctypes front end of the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
reference arm may import this module; nothing under ``rootsolve_b200/`` does.

``stodes_args`` restates the argument preprocessing of ``R/stodes.R:56-160``
(itol, Type, the ``nnz`` vector, default ``lrw`` and ``ngp``) independently of
the product's host mirror so that the two can be compared.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

DERIV = C.CFUNCTYPE(None, C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_double),
                    C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int))
PARMS_CB = C.CFUNCTYPE(None, C.POINTER(C.c_int), C.POINTER(C.c_double))
INIT = C.CFUNCTYPE(None, PARMS_CB)


class OrcArgs(C.Structure):
    _fields_ = [
        ("neq", C.c_int), ("time", C.c_double), ("func", C.c_void_p), ("ctol", C.c_double),
        ("atol", C.POINTER(C.c_double)), ("natol", C.c_int),
        ("rtol", C.POINTER(C.c_double)), ("nrtol", C.c_int),
        ("itol", C.c_int),
        ("nnzvec", C.POINTER(C.c_int)), ("nnnz", C.c_int),
        ("lrw", C.c_int), ("ngp", C.c_int), ("maxiter", C.c_int), ("posit", C.c_int),
        ("pos", C.POINTER(C.c_int)), ("ipos", C.c_int),
        ("nout", C.c_int),
        ("rpar", C.POINTER(C.c_double)), ("nrpar", C.c_int),
        ("ipar", C.POINTER(C.c_int)), ("nipar", C.c_int),
        ("type", C.c_int),
        ("ian_in", C.POINTER(C.c_int)), ("jan_in", C.POINTER(C.c_int)),
        ("imethod", C.c_int), ("quirk_first_step_transposed", C.c_int), ("quirk_unterminated_igp", C.c_int),
    ]


class OrcResult(C.Structure):
    _fields_ = [
        ("y", C.POINTER(C.c_double)), ("precis", C.POINTER(C.c_double)),
        ("niter", C.c_int), ("steady", C.c_int), ("dims", C.c_int * 3),
        ("ian", C.POINTER(C.c_int)), ("jan", C.POINTER(C.c_int)),
        ("igp", C.POINTER(C.c_int)), ("jgp", C.POINTER(C.c_int)),
        ("a0", C.POINTER(C.c_double)), ("beta0", C.POINTER(C.c_double)),
        ("t_struct", C.c_double), ("t_order", C.c_double), ("t_jac", C.c_double),
        ("t_factor", C.c_double), ("t_solve", C.c_double), ("t_total", C.c_double),
        ("status", C.c_int), ("msg", C.c_char * 256),
    ]


def build(force: bool = False) -> str:
    """Compile oracle/liboracle.so with the committed Makefile (gcc only)."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("stodes_oracle.c", "oracle_models.c", "stodes_oracle.h")]
    srcs.append(os.path.join(_HERE, "..", "rootsolve_b200", "models", "model_math.h"))
    stale = force or not os.path.exists(so) or any(
        os.path.getmtime(s) > os.path.getmtime(so) for s in srcs if os.path.exists(s))
    if stale:
        subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], check=True,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_stsparse.argtypes = [C.POINTER(OrcArgs), C.POINTER(OrcResult)]
        _LIB.orc_stsparse.restype = C.c_int
        _LIB.orc_perturb.argtypes = [C.POINTER(C.c_double)]
        _LIB.orc_perturb.restype = C.c_double
    return _LIB


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def sparse_structure(kind: str, nspec: int, dims, cyclic=None):
    """ian/jan (1-based) of sparse1d / sparse2d / sparse3d."""
    L = lib()
    dims = np.atleast_1d(np.asarray(dims, dtype=np.int32))
    cyc = np.zeros(len(dims), dtype=np.int32) if cyclic is None else np.asarray(cyclic, dtype=np.int32)
    n = int(nspec * np.prod(dims))
    cap = n * (2 * len(dims) + nspec)
    ian = np.zeros(n + 1, dtype=np.int32)
    jan = np.zeros(cap, dtype=np.int32)
    nnz = C.c_int(cap)
    if kind == "1D":
        rc = L.orc_sparse1d(n, nspec, int(dims[0]), int(cyc[0]), C.byref(nnz), _ip(ian), _ip(jan))
    elif kind == "2D":
        rc = L.orc_sparse2d(n, nspec, _ip(dims), _ip(cyc), C.byref(nnz), _ip(ian), _ip(jan))
    elif kind == "3D":
        rc = L.orc_sparse3d(n, nspec, _ip(dims), _ip(cyc), C.byref(nnz), _ip(ian), _ip(jan))
    else:
        raise ValueError(kind)
    if rc:
        raise RuntimeError(f"oracle structure generator failed rc={rc}")
    return ian, jan[: nnz.value].copy()


def mapped_structure(kind: str, nspec: int, dims, pres, cyclic=None, cap=None):
    """ian/jan (1-based, in the NEW numbering) of sparse2dmap / sparse3dmap; pres over the full grid."""
    L = lib()
    dims = np.atleast_1d(np.asarray(dims, dtype=np.int32))
    cyc = np.zeros(len(dims), dtype=np.int32) if cyclic is None else np.asarray(cyclic, dtype=np.int32)
    pres = np.ascontiguousarray(pres, dtype=np.int32).ravel()
    n = int((pres > 0).sum())
    cap = len(pres) * (2 * len(dims) + nspec) + 1 if cap is None else cap
    ian = np.zeros(n + 1, dtype=np.int32)
    jan = np.zeros(cap, dtype=np.int32)
    nnz = C.c_int(cap)
    fn = {"2Dmap": L.orc_sparse2dmap, "3Dmap": L.orc_sparse3dmap}[kind]
    rc = fn(n, nspec, _ip(dims), _ip(cyc), C.byref(nnz), _ip(ian), _ip(jan), _ip(pres), len(pres))
    if rc:
        raise RuntimeError(f"oracle structure generator failed rc={rc}")
    return ian, jan[: nnz.value].copy()


def jgroup(ian, jan, maxg=None):
    L = lib()
    n = len(ian) - 1
    maxg = n + 1 if maxg is None else maxg
    igp = np.zeros(maxg + 1, dtype=np.int32)
    jgp = np.zeros(n, dtype=np.int32)
    ngp = C.c_int(0)
    ian = np.ascontiguousarray(ian, dtype=np.int32)
    jan = np.ascontiguousarray(jan, dtype=np.int32)
    rc = L.orc_jgroup(n, _ip(ian), _ip(jan), maxg, C.byref(ngp), _ip(igp), _ip(jgp))
    if rc:
        raise RuntimeError(f"oracle jgroup failed rc={rc}")
    return ngp.value, igp[: ngp.value + 1].copy(), jgp


def stodes_args(n, sparsetype, nnz=None, ngp=None, lrw=None, atol=1e-8, rtol=1e-6):
    """R/stodes.R:56-160 restated: returns (Type, nnzvec, ngp, lrw, itol)."""
    atol = np.atleast_1d(atol)
    rtol = np.atleast_1d(rtol)
    itol = 1
    if len(atol) == n and len(rtol) == n:
        itol = 4
    elif len(atol) == n:
        itol = 2
    elif len(rtol) == n:
        itol = 3
    if ngp is None:
        ngp = n + 1
    if sparsetype == "1D":
        Type = 2
        nspec = int(nnz[0])
        v = [n * (2 + nspec) - 2 * nspec] + [int(q) for q in nnz]
        ngp = 3 * nspec + 1
        if v[3] == 1:
            v[0] += 2 * nspec
            ngp += 1
    elif sparsetype in ("2D", "2Dmap"):
        Type = 3 if sparsetype == "2D" else 30
        nspec = int(nnz[0])
        dimens = [int(q) for q in nnz[1:3]]
        nstate = n if Type == 3 else nspec * dimens[0] * dimens[1]      # R/stodes.R:116-119
        v = [nstate * (4 + nspec) - 2 * nspec * sum(dimens)] + [int(q) for q in nnz]
        dimmax = max(dimens)
        # "ngp < 4*nspec+1" in the reference is a comparison, not an assignment
        for slot in (4, 5):
            if v[slot] == 1:
                v[0] += 2 * dimmax * nspec
                ngp += 1
    elif sparsetype in ("3D", "3Dmap"):
        Type = 4 if sparsetype == "3D" else 40
        nspec = int(nnz[0])
        dimens = [int(q) for q in nnz[1:4]]
        nstate = n if Type == 4 else nspec * dimens[0] * dimens[1] * dimens[2]   # R/stodes.R:138-141
        v = [nstate * (6 + nspec) - 3 * nspec * sum(dimens)] + [int(q) for q in nnz]
        if v[5] == 1:
            v[0] += 2 * dimens[1] * dimens[2] * nspec
            ngp += 1
        if v[6] == 1:
            v[0] += 2 * dimens[0] * dimens[2] * nspec
            ngp += 1
        if v[7] == 1:
            v[0] += 2 * dimens[0] * dimens[1] * nspec
            ngp += 1
    elif sparsetype in ("sparseint", "sparsereturn", "sparse"):
        Type = 1
        v = [n * n if nnz is None else int(nnz)]
    else:
        raise ValueError("cannot run stodes: sparsetype not known ")
    if lrw is None:
        lrw = 3 * n + 4 * v[0]
    return Type, np.asarray(v, dtype=np.int32), int(ngp), int(lrw), itol


_KEEP = []  # keep ctypes callbacks alive


def model(name: str, parms, forcs=None):
    """Address of a bundled host model after running its initialiser(s) with parms (and forcs)."""
    L = lib()
    p = np.ascontiguousarray(parms, dtype=np.float64)

    def steadyparms(N, dst):
        if N[0] != len(p):
            raise ValueError("Confusion over the length of parms.")
        for i in range(N[0]):
            dst[i] = p[i]

    cb = PARMS_CB(steadyparms)
    _KEEP.append(cb)
    getattr(L, "init" + name)(cb)
    if forcs is not None:       # the forcing initialiser, src/call_stsparse.c:242-251
        f = np.ascontiguousarray(forcs, dtype=np.float64)

        bad = []

        def steadyforcs(N, dst):
            if N[0] != len(f):      # an exception cannot cross the C frame: note it, raise after the call
                bad.append(N[0])
                return
            for i in range(N[0]):
                dst[i] = f[i]

        cbf = PARMS_CB(steadyforcs)
        _KEEP.append(cbf)
        getattr(L, "forc" + name)(cbf)
        if bad:
            raise ValueError("Confusion over the length of forcs.")
    return C.cast(getattr(L, name), C.c_void_p)


def stsparse(y, func, *, sparsetype, nnz, time=0.0, atol=1e-8, rtol=1e-6, ctol=1e-8, lrw=None,
             ngp=None, positive=False, maxiter=100, ipar=None, rpar=None, nout=0,
             quirk_first_step_transposed=True, quirk_unterminated_igp=True, want_jacobian=False, nnz_cap=None):
    """Run the oracle's whole solve; returns a dict shaped like stodes' result.  nnz_cap overrides the nonzero
    estimate R computes (tests only: the R-level stodes gives the user no such handle for grid types)."""
    L = lib()
    y = np.ascontiguousarray(y, dtype=np.float64).ravel()
    n = len(y)
    ret_only = sparsetype == "sparsereturn"
    Type, nnzvec, maxg, lrw, itol = stodes_args(n, sparsetype, nnz, ngp, lrw, atol, rtol)
    imethod = 1
    if ret_only:
        imethod, Type = 10, -Type
    if nnz_cap is not None:
        nnzvec[0] = nnz_cap
    atol_a = np.atleast_1d(np.asarray(atol, dtype=np.float64))
    rtol_a = np.atleast_1d(np.asarray(rtol, dtype=np.float64))
    if isinstance(positive, (bool, np.bool_)):
        posit, pos = int(positive), np.asarray([int(positive)], dtype=np.int32)
    else:
        posit, pos = 0, np.asarray(positive, dtype=np.int32)
    ipar_a = np.zeros(0, dtype=np.int32) if ipar is None else np.asarray(ipar, dtype=np.int32)
    rpar_a = np.zeros(0, dtype=np.float64) if rpar is None else np.asarray(rpar, dtype=np.float64)
    cap = int(nnzvec[0])
    yout = np.zeros(n + nout)
    yout[:n] = y
    precis = np.zeros(maxiter)
    ian = np.zeros(n + 1, dtype=np.int32)
    jan = np.zeros(cap, dtype=np.int32)
    igp = np.zeros(maxg + 1, dtype=np.int32)
    jgp = np.zeros(n, dtype=np.int32)
    a0 = np.zeros(cap if want_jacobian else 1)
    beta0 = np.zeros(n if want_jacobian else 1)
    A = OrcArgs(neq=n, time=time, func=func, ctol=ctol, atol=_dp(atol_a), natol=len(atol_a),
                rtol=_dp(rtol_a), nrtol=len(rtol_a), itol=itol, nnzvec=_ip(nnzvec), nnnz=len(nnzvec),
                lrw=lrw, ngp=maxg, maxiter=maxiter, posit=posit, pos=_ip(pos), ipos=len(pos), nout=nout,
                rpar=_dp(rpar_a), nrpar=len(rpar_a), ipar=_ip(ipar_a), nipar=len(ipar_a), type=Type,
                ian_in=None, jan_in=None, imethod=imethod,
                quirk_first_step_transposed=int(quirk_first_step_transposed),
                quirk_unterminated_igp=int(quirk_unterminated_igp))
    R = OrcResult(y=_dp(yout), precis=_dp(precis), ian=_ip(ian), jan=_ip(jan), igp=_ip(igp), jgp=_ip(jgp),
                  a0=_dp(a0) if want_jacobian else None, beta0=_dp(beta0) if want_jacobian else None)
    rc = L.orc_stsparse(C.byref(A), C.byref(R))
    if rc:
        raise RuntimeError(f"oracle: {R.msg.decode()} (code {rc})")
    nz, g = R.dims[0], R.dims[1]
    out = dict(y=yout[:n].copy(), out=yout[n:].copy(), precis=precis[: R.niter].copy(), steady=bool(R.steady),
               dims=dict(nnz=nz, ngp=g, lrw=R.dims[2]), ian=ian, jan=jan[:nz].copy(),
               igp=igp[: g + 1].copy(), jgp=jgp,
               timing=dict(struct=R.t_struct, order=R.t_order, jac=R.t_jac, factor=R.t_factor,
                           solve=R.t_solve, total=R.t_total))
    if want_jacobian:
        out["a0"], out["beta0"] = a0[:nz].copy(), beta0
    return out
