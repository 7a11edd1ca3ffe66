"""python scratch/setup_pinned.py : rs_stsparse on C5 256^3 with PINNED host buffers (bench.py's GpuCase), four calls;
prints the library's own set-up / Newton times of each.  Run once plain and once with RS_DEBUG_SETUP=1 (phase times on
stderr; the debug switch also keeps the state upload on the solver's stream)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from rootsolve_b200 import _capi
host = importlib.import_module("rootsolve_b200.stodes")
torch.cuda.set_device(0)
L = _capi.lib()
G = bench.GpuCase(bench.case("3D", 256), host, _capi, torch)
import ctypes as C
for it in range(4):
    R = G.result(True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    _capi.check(L.rs_stsparse(C.byref(G.P.args), C.byref(R)))
    torch.cuda.synchronize()
    sys.stderr.write("call %d: wall %.2f ms  setup %.2f ms  newton %.2f ms\n" % (it, 1e3 * (time.perf_counter() - t0), 1e3 * R.t_setup_s, 1e3 * R.t_newton_s))
