"""Multi-GPU plumbing: one process per GPU, slabs over the slowest grid index.

``torch.distributed`` is used only to ship the 128-byte NCCL unique id from rank 0
to the other ranks; all data-path communication (halo planes, dot-product and
convergence all-reduces) happens inside the CUDA library over its own NCCL
communicator (``rootsolve_b200/csrc/comm.cu``).

The index helpers below are the host-side statement of the partition the
library uses (``rs_slab_bounds`` + species-major layout); they are
pure Python so that the world_size-2 ``gloo`` tests can exercise them on CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np


def _slab_bound(d1: int, nranks: int, r: int) -> int:
    if r <= 0:
        return 0
    if r >= nranks:
        return d1
    base, rem = divmod(d1, nranks)
    A = 1
    while A < 16 and d1 >= nranks * 4 * (2 * A):
        A *= 2
    return (r * base + min(r, rem)) // A * A


def slab_bounds(d1: int, nranks: int, rank: int):
    """Planes [j0, j1) of the slowest grid dimension owned by ``rank`` (same rule as rs_slab_bounds): an even split
    whose interior bounds are moved down to a multiple of 2..16 planes while every rank keeps enough of them -- whole
    2^k aggregates per rank are what lets the finest multigrid levels span the ranks."""
    return _slab_bound(d1, nranks, rank), _slab_bound(d1, nranks, rank + 1)


def owned_ranges(nspec: int, dims, nranks: int, rank: int):
    """Global index ranges [b, e) a rank owns: one per species (dims = rev(dimens), dims[0] slowest)."""
    ncell = int(np.prod(dims))
    plane = ncell // dims[0]
    j0, j1 = slab_bounds(dims[0], nranks, rank)
    return [(s * ncell + j0 * plane, s * ncell + j1 * plane) for s in range(nspec)]


def halo_ranges(nspec: int, dims, nranks: int, rank: int):
    """(recv_from_lower, recv_from_upper, send_to_lower, send_to_upper) index ranges per species; None at the ends."""
    ncell = int(np.prod(dims))
    plane = ncell // dims[0]
    j0, j1 = slab_bounds(dims[0], nranks, rank)
    out = []
    for s in range(nspec):
        base = s * ncell
        lo = rank > 0
        hi = rank < nranks - 1
        out.append(dict(
            recv_lo=(base + (j0 - 1) * plane, base + j0 * plane) if lo else None,
            send_lo=(base + j0 * plane, base + (j0 + 1) * plane) if lo else None,
            recv_hi=(base + j1 * plane, base + (j1 + 1) * plane) if hi else None,
            send_hi=(base + (j1 - 1) * plane, base + j1 * plane) if hi else None))
    return out


def exchange_halos(vec, nspec, dims, dist):
    """Reference halo exchange on a host (numpy) vector through torch.distributed point-to-point calls --
    what comm_halo_exchange does with ncclSend/ncclRecv on the device.  Used by the CPU tests."""
    import torch
    rank, world = dist.get_rank(), dist.get_world_size()
    ops, bufs = [], []
    for h in halo_ranges(nspec, dims, world, rank):
        for key_s, key_r, peer in (("send_lo", "recv_lo", rank - 1), ("send_hi", "recv_hi", rank + 1)):
            if h[key_s] is None:
                continue
            sb = torch.from_numpy(np.ascontiguousarray(vec[h[key_s][0]:h[key_s][1]]))
            rb = torch.empty(h[key_r][1] - h[key_r][0], dtype=torch.float64)
            ops += [dist.P2POp(dist.isend, sb, peer), dist.P2POp(dist.irecv, rb, peer)]
            bufs.append((h[key_r], rb))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    for (b, e), rb in bufs:
        vec[b:e] = rb.numpy()
    return vec


def init_comm(dist) -> None:
    """Creates the library's NCCL communicator; call once per process after init_process_group
    and after selecting this rank's GPU (torch.cuda.set_device)."""
    import torch
    from . import _capi
    L = _capi.lib()
    rank, world = dist.get_rank(), dist.get_world_size()
    if world == 1:
        return
    buf = C.create_string_buffer(128)
    if rank == 0:
        _capi.check(L.rs_comm_unique_id(buf))
    t = torch.tensor(list(buf.raw), dtype=torch.uint8)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.broadcast(t, src=0)
    raw = bytes(t.cpu().tolist())
    _capi.check(L.rs_comm_init(world, rank, raw))


def finalize_comm() -> None:
    from . import _capi
    _capi.lib().rs_comm_finalize()
