import os, sys
os.environ["RS_DEBUG_SETUP"] = "1"
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, ctypes as C, importlib, time
import helpers as H
from rootsolve_b200 import _capi
host = importlib.import_module("rootsolve_b200.stodes")
n = 256
y0 = np.full(n**3, 10.0)
P = host.pack_args(y0, 0.0, "diff3d", H.P_DIFF3D, 1e-10, 1e-10, 1e-10, "3D", False, [1, n, n, n, 0, 0, 0], None, None, None, True, 100, None, "rs_models", H.P_DIFF3D, None, [1, n, n, n, 0], 0, "yale", None)
L = _capi.lib()
for rep in range(int(os.environ.get('REPS','3'))):
    h = C.c_void_p()
    t0 = time.perf_counter()
    _capi.check(L.rs_solver_create(C.byref(P.args), C.byref(h)))
    t1 = time.perf_counter()
    L.rs_solver_destroy(h)
    t2 = time.perf_counter()
    print("create %.1f ms  destroy %.1f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3), file=sys.stderr)
