import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
import rootsolve_b200 as rs
import helpers as H
n = 256
y0 = np.full(n**3, 10.0)
for rep in range(3):
    out = rs.steady_3D(y0, func="diff3d", parms=H.P_DIFF3D, dimens=[n]*3, positive=True, ipar=[1, n, n, n, 0], maxiter=1, **H.TOL) if False else None
import ctypes as C
from rootsolve_b200 import _capi
import importlib
host = importlib.import_module("rootsolve_b200.stodes")
P = host.pack_args(y0, 0.0, "diff3d", H.P_DIFF3D, 1e-10, 1e-10, 1e-10, "3D", False, [1, n, n, n, 0, 0, 0], None, None, None, True, 100, None, "rs_models", H.P_DIFF3D, None, [1, n, n, n, 0], 0, "yale", None)
L = _capi.lib()
for rep in range(3):
    h = C.c_void_p()
    t0 = time.perf_counter()
    _capi.check(L.rs_solver_create(C.byref(P.args), C.byref(h)))
    t1 = time.perf_counter()
    L.rs_solver_destroy(h)
    print("create %.1f ms" % ((t1 - t0) * 1e3))
