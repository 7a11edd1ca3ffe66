import sys, os
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, ctypes as C
import rootsolve_b200 as rs
from rootsolve_b200 import _capi
import helpers as H
d = int(sys.argv[1]) if len(sys.argv) > 1 else 128
nc = d**3
ipar = [2, d, d, d, 0]
y0 = np.concatenate([np.full(nc, 10.0), np.full(nc, 1.0)])
out = rs.steady_3D(y0, func="rd3d", parms=H.P_RD3D, nspec=2, dimens=[d, d, d], positive=True, ipar=ipar, verbose=True, **H.TOL)
A = out.attributes
print(A["steady"], A["precis"], A["krylov_iters"], A["dims"])
y = out["y"]; print("y range", y.min(), y.max(), np.isnan(y).sum())
L = _capi.lib()
f = np.zeros(2*nc); p = np.asarray(H.P_RD3D); ip = np.asarray(ipar, dtype=np.int32)
_capi.check(L.rs_residual(_capi.model_symbol("rd3d"), H.dp(p), len(p), 2*nc, 0.0, H.dp(y), None, 0, H.ip(ip), len(ip), H.dp(f)))
print("gpu residual max", np.abs(f).max(), "argmax", np.abs(f).argmax())
