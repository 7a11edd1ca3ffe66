import sys, numpy as np
sys.path.insert(0, "scratch")
from mg_proto import *
import mg_proto

S = 64
A3 = grid_matrix([S, S, S], 0.025); A3n = grid_matrix([S, S, S], 1e-4, dirichlet=False)
A2 = grid_matrix([512, 512], 0.0, dirichlet=False, rvar=np.random.default_rng(3).uniform(0, 0.02, 512 * 512))
rng = np.random.default_rng(1)
for name, A, dims in [("C5-64", A3, [S, S, S]), ("3Dneumann-64", A3n, [S, S, S]), ("2Dneumann-512 rvar", A2, [512, 512])]:
    b = rng.standard_normal(A.shape[0])
    for ratio in (4.0, 6.0, 8.0, 12.0, 20.0):
        # patch lmin ratio
        src = mg_proto.MG.smooth
        class M2(MG):
            def smooth(self, lev, x, b_, first_zero, ratio=ratio):
                A_, dinv = lev["A"], lev["dinv"]
                lmax = lev["lmax"]; lmin = lmax / ratio
                theta = 0.5 * (lmax + lmin); delta = 0.5 * (lmax - lmin)
                sigma = theta / delta; rho = 1.0 / sigma
                r = b_ - A_ @ x if not first_zero else b_.copy()
                d = dinv * r / theta
                x = x + d
                for _ in range(self.cheb - 1):
                    r = r - A_ @ d
                    rho_new = 1.0 / (2 * sigma - rho)
                    d = rho_new * rho * d + 2 * rho_new / delta * (dinv * r)
                    rho = rho_new
                    x = x + d
                return x
        for fix in (0.5,):
            for cheb in (1, 2, 3):
                M = M2(A, dims, fix=fix, cheb=cheb)
                x, it = pcg(A, b, M)
                print(f"{name:20s} ratio={ratio:5.1f} fix={fix} cheb={cheb} its={it}", flush=True)
