"""Where the set-up time of config C2 (steady.1D, 5 x 200000) goes: RS_DEBUG_SETUP=1 python scratch/c2_setup.py"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rootsolve_b200 as rs
nb, ns = 200000, 5
for rep in range(2):
    t0 = time.time()
    out = rs.steady_1D(np.full(ns * nb, 1.0), nspec=ns, method="stodes", func="rt1d", parms=[1.0, 1.0, 1.0, 0.5, 0.5, 1.0, 0.1],
                       positive=True, ipar=[ns, nb, 1, 1, 0], atol=1e-10, rtol=1e-10, ctol=1e-10)
    print("wall", time.time() - t0, out.attributes["timing"], out.attributes["krylov_iters"], out.attributes["precond"], flush=True)
