import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
import rootsolve_b200 as rs
from oracle import oracle as O
import helpers as H
for n in (50, 100, 200):
    y0 = np.full(n*n, 10.0)
    kw = dict(func="diff2d", parms=H.P_DIFF2D_MONOD, dimens=[n, n], positive=True, ipar=[1, n, n, 1, 1], **H.TOL)
    rs.steady_2D(y0, **kw)
    t0 = time.perf_counter(); out = rs.steady_2D(y0, **kw); tg = time.perf_counter() - t0
    t0 = time.perf_counter()
    ref = O.stsparse(y0, O.model("diff2d", H.P_DIFF2D_MONOD), sparsetype="2D", nnz=[1, n, n, 0, 0], ipar=[1, n, n, 1, 1], lrw=40*n*n*12, positive=True, **H.TOL)
    tc = time.perf_counter() - t0
    A = out.attributes
    print(f"2D {n}x{n} Monod: GPU {tg*1e3:.1f} ms (setup {A['timing']['setup_s']*1e3:.1f}, newton {A['timing']['newton_s']*1e3:.1f}; {len(A['precis'])} newton, {A['krylov_iters']} krylov, {A['kernel_launches']} launches)  CPU oracle {tc*1e3:.1f} ms ({len(ref['precis'])} newton)")
