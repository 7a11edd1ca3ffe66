import sys, os
sys.path.insert(0, '.')
sys.path.insert(0, 'tests')
import numpy as np
import rootsolve_b200 as rs
import helpers as H
y0 = np.full(5*2000, 1.0)
try:
    out = rs.steady_1D(y0, nspec=5, method="stodes", func="rt1d", parms=H.P_RT1D, positive=True, ipar=[5,2000,1,1,0], verbose=True, **H.TOL)
    print(out.attributes)
except Exception as e:
    print("ERR", e)
