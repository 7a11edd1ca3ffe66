"""ctypes binding of the C ABI declared in ``include/rootsolve_b200.h``.

The library is the product: if ``librootsolve_b200.so`` is missing the import
of this module raises -- there is no Python or CPU fallback behind it.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_SOLVER = os.path.join(HERE, "librootsolve_b200.so")
LIB_MODELS = os.path.join(HERE, "librs_models.so")

RS_OK = 0
RS_TYPE_USER, RS_TYPE_1D, RS_TYPE_2D, RS_TYPE_3D = 0, 2, 3, 4
RS_TYPE_2DMAP, RS_TYPE_3DMAP = 30, 40
RS_PRECOND_NONE, RS_PRECOND_JACOBI, RS_PRECOND_BLOCK_JACOBI, RS_PRECOND_MG, RS_PRECOND_BTD, RS_PRECOND_AUTO = 0, 1, 2, 3, 4, -1
ERROR_NAMES = {-1: "RS_EINVAL", -2: "RS_ENODEVICE", -3: "RS_ENOMEM", -4: "RS_ESTRUCT", -5: "RS_EGROUP",
               -6: "RS_EDUPLICATE", -7: "RS_ESINGULAR", -8: "RS_EMODEL", -9: "RS_ECOMM", -10: "RS_ETOOBIG"}

c_int_p = C.POINTER(C.c_int)
c_double_p = C.POINTER(C.c_double)


class KrylovControl(C.Structure):
    _fields_ = [("precond", C.c_int), ("max_krylov", C.c_int), ("lin_rtol", C.c_double),
                ("lin_ewt_factor", C.c_double), ("check_every", C.c_int)]


class StsparseArgs(C.Structure):
    _fields_ = [
        ("y", c_double_p), ("neq", C.c_int), ("time", C.c_double),
        ("func", C.c_void_p),
        ("initpar", c_double_p), ("ninitpar", C.c_int),
        ("forcs", c_double_p), ("nforcs", C.c_int),
        ("ctol", C.c_double), ("atol", c_double_p), ("natol", C.c_int), ("rtol", c_double_p), ("nrtol", C.c_int),
        ("itol", C.c_int), ("verbose", C.c_int),
        ("nnz", c_int_p), ("nnnz", C.c_int),
        ("lrw", C.c_int), ("ngp", C.c_int), ("maxiter", C.c_int), ("Pos", C.c_int),
        ("positive", c_int_p), ("npositive", C.c_int),
        ("nout", C.c_int),
        ("rpar", c_double_p), ("nrpar", C.c_int),
        ("ipar", c_int_p), ("nipar", C.c_int),
        ("Type", C.c_int),
        ("ian", c_int_p), ("jan", c_int_p),
        ("imethod", C.c_int),
        ("control", C.POINTER(KrylovControl)),
    ]


class StsparseResult(C.Structure):
    _fields_ = [
        ("y", c_double_p), ("precis", c_double_p), ("niter", C.c_int), ("steady", C.c_int * 4),
        ("ian", c_int_p), ("jan", c_int_p), ("igp", c_int_p), ("jgp", c_int_p),
        ("krylov_iters", C.c_int), ("kernel_launches", C.c_int),
        ("t_setup_s", C.c_double), ("t_newton_s", C.c_double), ("spmv_kernel", C.c_int),
        ("precond", C.c_int), ("krylov_method", C.c_int),
    ]


class RootSolveError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{ERROR_NAMES.get(code, code)}: {msg}")
        self.code = code


_lib = None
_model_libs = {}


def lib():
    """The solver library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_SOLVER):
            raise ImportError(
                f"{LIB_SOLVER} is missing: build it with `python -m rootsolve_b200.build` "
                "(nvcc, sm_100a). rootsolve_b200 has no CPU fallback.")
        L = C.CDLL(LIB_SOLVER, mode=C.RTLD_GLOBAL)
        L.rs_last_error.restype = C.c_char_p
        L.rs_stsparse.argtypes = [C.POINTER(StsparseArgs), C.POINTER(StsparseResult)]
        L.rs_solver_create.argtypes = [C.POINTER(StsparseArgs), C.POINTER(C.c_void_p)]
        L.rs_solver_set_state.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.rs_solver_run.argtypes = [C.c_void_p, C.POINTER(StsparseResult)]
        L.rs_solver_state_device.argtypes = [C.c_void_p]
        L.rs_solver_state_device.restype = C.c_void_p
        L.rs_solver_get_state.argtypes = [C.c_void_p, c_double_p]
        L.rs_solver_set_profiling.argtypes = [C.c_void_p, C.c_int]
        L.rs_solver_kernel_stats.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), c_double_p,
                                             C.POINTER(C.c_longlong), c_double_p]
        L.rs_solver_destroy.argtypes = [C.c_void_p]
        L.rs_solver_destroy.restype = None
        L.rs_sparse_structure.argtypes = [C.c_int, C.c_int, C.c_int, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p]
        L.rs_sparse_structure_map.argtypes = [C.c_int, C.c_int, C.c_int, c_int_p, c_int_p, c_int_p, C.c_int, c_int_p,
                                              c_int_p, c_int_p]
        L.rs_jgroup.argtypes = [C.c_int, c_int_p, c_int_p, C.c_int, c_int_p, c_int_p, c_int_p, c_int_p]
        L.rs_fd_jacobian.argtypes = [C.c_void_p, c_double_p, C.c_int, C.c_int, C.c_double, c_double_p, c_int_p, c_int_p,
                                     c_int_p, c_int_p, C.c_int, c_double_p, C.c_int, c_int_p, C.c_int, c_double_p,
                                     c_double_p]
        L.rs_residual.argtypes = [C.c_void_p, c_double_p, C.c_int, C.c_int, C.c_double, c_double_p, c_double_p, C.c_int,
                                  c_int_p, C.c_int, c_double_p]
        L.rs_linsolve.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, C.c_int, C.c_int,
                                  C.POINTER(KrylovControl), c_double_p, c_int_p, c_double_p]
        L.rs_linsolve_grid.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_int_p,
                                       C.POINTER(KrylovControl), c_double_p, c_int_p, c_double_p, c_int_p]
        L.rs_spmv.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_double_p]
        L.rs_errset.argtypes = [C.c_int, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p]
        L.rs_convergence_test.argtypes = [C.c_int, c_double_p, c_double_p, c_double_p]
        L.rs_convergence_test_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rs_release_cached_memory.restype = C.c_longlong
        L.rs_comm_unique_id.argtypes = [C.c_char_p]
        L.rs_comm_init.argtypes = [C.c_int, C.c_int, C.c_char_p]
        L.rs_comm_finalize.restype = None
        L.rs_slab_bounds.argtypes = [C.c_int, C.c_int, C.c_int, c_int_p, c_int_p]
        L.rs_slab_bounds.restype = None
        _lib = L
    return _lib


def check(rc):
    if rc != RS_OK:
        raise RootSolveError(rc, lib().rs_last_error().decode(errors="replace"))


def model_symbol(func: str, dllname: str | None = None) -> int:
    """Address of ``rs_model_<func>()``'s descriptor in the model library ``dllname``.

    Mirrors ``getNativeSymbolInfo(func, PACKAGE = dllname)$address`` (R/stodes.R:212-220).
    ``dllname`` is a path, or a bare name looked up next to this package; default: the
    bundled ``rs_models`` library.
    """
    path = LIB_MODELS if dllname in (None, "rs_models") else dllname
    if not os.path.exists(path):
        cand = os.path.join(HERE, f"lib{dllname}.so")
        if os.path.exists(cand):
            path = cand
        else:
            raise ImportError(f"model library {path!r} not found (no CPU fallback; build with "
                              "`python -m rootsolve_b200.build`)")
    L = _model_libs.get(path)
    if L is None:
        # RTLD_LOCAL: two model libraries may instantiate kernels of the same name (copies of one model); each must
        # keep its own
        L = _model_libs[path] = C.CDLL(path, mode=C.RTLD_LOCAL)
    try:
        getter = getattr(L, f"rs_model_{func}")
    except AttributeError as e:
        raise ValueError(f"cannot calculate steady-state: dyn function not loaded: {func}") from e
    getter.restype = C.c_void_p
    return getter()
