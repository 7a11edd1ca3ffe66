"""World-size-2 tests of the slab partition on CPU (gloo): the host statement of what every rank owns and
exchanges (rootsolve_b200/parallel.py, the same rule as rs_slab_bounds / comm_halo_exchange) must let each rank
compute, from its slab plus one halo plane per side, exactly the residual the unpartitioned model gives."""
import ctypes as C
import os
import socket

import numpy as np
import pytest

import helpers as H


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, case, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import sys
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        sys.path.insert(0, root)
        sys.path.insert(0, os.path.join(root, "tests"))
        from oracle import oracle as O
        from rootsolve_b200 import parallel as P
        model, parms, nspec, dims, ipar = case
        n = nspec * int(np.prod(dims))
        rng = np.random.default_rng(20260101)            # same vector on every rank
        y_global = rng.uniform(0.5, 10.0, n)
        func = C.cast(O.model(model, parms), O.DERIV)
        ipfull = np.concatenate([[0, 0, 3 + len(ipar)], ipar]).astype(np.int32)

        def residual(y):
            f = np.zeros(n)
            neq, t = C.c_int(n), C.c_double(0.0)
            func(C.byref(neq), C.byref(t), H.dp(y), H.dp(f), H.dp(np.zeros(1)), H.ip(ipfull))
            return f
        f_global = residual(y_global.copy())
        # this rank only knows its own planes: everything else is poisoned, then halos are exchanged
        mine = P.owned_ranges(nspec, dims, world, rank)
        y_local = np.full(n, np.nan)
        for b, e in mine:
            y_local[b:e] = y_global[b:e]
        P.exchange_halos(y_local, nspec, dims, dist)
        f_local = residual(np.nan_to_num(y_local, nan=1e300))   # poison must not reach owned rows
        ok = all(np.array_equal(f_local[b:e], f_global[b:e]) for b, e in mine)
        # the partition covers every unknown exactly once
        import torch
        cover = np.zeros(n)
        for b, e in mine:
            cover[b:e] += 1
        t = torch.from_numpy(cover)
        dist.all_reduce(t)
        ok = ok and bool(np.all(t.numpy() == 1))
        # scalars: a sum all-reduce of the per-slab partial sums reproduces the global sum of |f|
        part = torch.tensor([sum(np.abs(f_local[b:e]).sum() for b, e in mine)], dtype=torch.float64)
        dist.all_reduce(part)
        ok = ok and abs(part.item() - np.abs(f_global).sum()) <= 1e-12 * np.abs(f_global).sum()
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


CASES = [
    ("diff3d", H.P_DIFF3D, 1, (7, 5, 6), [1, 7, 5, 6, 0]),
    ("rd3d", H.P_RD3D, 2, (6, 4, 5), [2, 6, 4, 5, 0]),
    ("diff2d", H.P_DIFF2D_MONOD, 1, (9, 8), [1, 9, 8, 1, 1]),
    ("rt1d", H.P_RT1D, 5, (31,), [5, 31, 1, 1, 0]),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_slab_partition_and_halo_exchange_world2(oracle, case):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, case, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    results = dict(q.get(timeout=5) for _ in range(2))
    assert results == {0: True, 1: True}


def test_partition_helpers_match_the_library(built):
    from rootsolve_b200 import parallel as P
    for d1, nr in [(256, 8), (10, 3), (7, 7), (128, 2)]:
        for r in range(nr):
            j0, j1 = C.c_int(0), C.c_int(0)
            built.rs_slab_bounds(d1, nr, r, C.byref(j0), C.byref(j1))
            assert P.slab_bounds(d1, nr, r) == (j0.value, j1.value)
    rng = P.owned_ranges(2, (8, 3, 4), 4, 1)
    assert rng == [(2 * 12, 4 * 12), (96 + 2 * 12, 96 + 4 * 12)]
    h = P.halo_ranges(1, (8, 3, 4), 4, 0)[0]
    assert h["recv_lo"] is None and h["send_hi"] == (12, 24) and h["recv_hi"] == (24, 36)
