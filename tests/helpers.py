"""Shared workload definitions for the tests (BASELINE.json configs at test sizes)."""
import ctypes as C

import numpy as np

TOL = dict(atol=1e-10, rtol=1e-10, ctol=1e-10)

# parameter vectors of the bundled models (rootsolve_b200/models/model_math.h)
P_DIFF2D_LIN = [1.0, 1.0, 1.0, 1.0, 0.025, 1.0, 1.0]
P_DIFF2D_MONOD = [1.0, 1.0, 1.0, 1.0, 0.25, 1.0, 1.0]
P_DIFF3D = [1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 0.025, 1.0]
# D1 D2 dx dy dz k m BNDu BNDv -- m > k*BNDu keeps dv/dv < 0 everywhere (a decaying, not a growing, species v)
P_RD3D = [1.0, 0.5, 1.0, 1.0, 1.0, 0.01, 0.2, 10.0, 1.0, 0.0]
P_RT1D = [1.0, 1.0, 1.0, 0.5, 0.5, 1.0, 0.1]


def ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def within_tol(y, yref, atol=1e-10, rtol=1e-10, factor=1.0):
    """The parity bound of BASELINE.json: |y - yref| <= atol + rtol*|yref| (times factor)."""
    err = np.abs(np.asarray(y) - np.asarray(yref))
    bound = factor * (atol + rtol * np.abs(yref))
    return bool(np.all(err <= bound)), float(np.max(err / bound))
