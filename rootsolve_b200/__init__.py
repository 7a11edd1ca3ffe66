"""rootsolve_b200 -- rootSolve's sparse-Newton steady-state path (stodes behind
steady.1D / steady.2D / steady.3D) on NVIDIA B200.

The package is a thin host mirror of the R API over a C-ABI CUDA library
(``include/rootsolve_b200.h``); see DESIGN.md.  There is no CPU fallback.
"""
from .stodes import stodes, steady, steady_1D, steady_2D, steady_3D, SteadyResult  # noqa: F401
from ._capi import RootSolveError  # noqa: F401


def release_cached_memory():
    """Returns the device work arrays cached between solves to the CUDA driver; the number of bytes released."""
    from . import _capi
    return int(_capi.lib().rs_release_cached_memory())


__all__ = ["stodes", "steady", "steady_1D", "steady_2D", "steady_3D", "SteadyResult", "RootSolveError",
           "release_cached_memory"]
