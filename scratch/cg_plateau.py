"""The evidence behind pcg_solve's plateau rule: Jacobi-PCG on a bare 500-box 1-D Laplacian (imposed value
left, zero gradient right; condition number 4e5), with the driver's check period of 4 steps.

  restart after 40 steps without a 10 % gain (the old rule):  not converged after 5000 steps, 71 restarts
  no restarts:                                                converged to 1e-8 in 500 steps (= N: finite termination)

numpy/scipy only; nothing of the product runs here.
"""
import numpy as np
import scipy.sparse as sp

N, D, dx = 500, 1.0, 0.2
main = -2 * np.ones(N) * D / dx ** 2
main[-1] = -D / dx ** 2
A = -sp.diags([np.ones(N - 1) * D / dx ** 2, main, np.ones(N - 1) * D / dx ** 2], [-1, 0, 1]).tocsr()
b = np.random.default_rng(0).standard_normal(N)


def pcg(restart_after, maxit=5000, tol=1e-8):
    d = A.diagonal()
    x = np.zeros(N)
    r = b.copy()
    z = r / d
    p = z.copy()
    rz = r @ z
    bn = np.linalg.norm(b)
    best, since, restarts = bn, 0, 0
    for it in range(1, maxit + 1):
        q = A @ p
        a = rz / (p @ q)
        x += a * p
        r -= a * q
        z = r / d
        rz2 = r @ z
        p = z + (rz2 / rz) * p
        rz = rz2
        if it % 4:
            continue
        rn = np.linalg.norm(r)
        if rn <= tol * bn:
            return it, restarts, rn / bn
        if rn < 0.9 * best:
            best, since = rn, 0
        else:
            since += 1
        if restart_after and since * 4 >= restart_after:
            r = b - A @ x
            z = r / d
            p = z.copy()
            rz = r @ z
            best, since, restarts = np.linalg.norm(r), 0, restarts + 1
    return maxit, restarts, np.linalg.norm(b - A @ x) / bn


if __name__ == "__main__":
    print("condition number      %.3g" % np.linalg.cond(A.toarray()))
    print("restart after 40      steps %d  restarts %d  relres %.2e" % pcg(40))
    print("restart after N = 500 steps %d  restarts %d  relres %.2e" % pcg(N))
    print("no restarts           steps %d  restarts %d  relres %.2e" % pcg(0))
