"""RS_DEBUG_SETUP=1 python scratch/setup_breakdown.py : set-up phases of C1 (50x50) and C5 (256^3), second call of each"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rootsolve_b200 as rs
TOL = dict(atol=1e-10, rtol=1e-10, ctol=1e-10)
for name, call in [("C1 50x50", lambda: rs.steady_2D(np.full(2500, 10.0), func="diff2d", parms=[1, 1, 1, 1, 0.25, 1, 1], dimens=[50, 50],
                                                     positive=True, ipar=[1, 50, 50, 1, 1], **TOL)),
                   ("C5 256^3", lambda: rs.steady_3D(np.full(256 ** 3, 10.0), func="diff3d", parms=[1, 1, 1, 1, 1, 1, 0.025, 1],
                                                     dimens=[256] * 3, positive=True, ipar=[1, 256, 256, 256, 0], **TOL))]:
    call()
    sys.stderr.write("---- %s\n" % name)
    t0 = time.time()
    out = call()
    sys.stderr.write("wall %.4f s  %s  newton %d krylov %d\n" % (time.time() - t0, out.attributes["timing"], len(out.attributes["precis"]),
                                                                 out.attributes["krylov_iters"]))
