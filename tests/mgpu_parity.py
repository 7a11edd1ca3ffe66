"""Multi-GPU parity driver (launched by torchrun, one rank per GPU; see tests/test_multi_gpu.py):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/mgpu_parity.py

Every rank first solves each case on its own GPU alone (no communicator), then the ranks solve it together
(slab-sharded, NCCL halos + all-reduces).  The sharded result must agree with the single-GPU result within the
reference's atol + rtol*|y| and with identical nnz / ngp / Newton iteration count."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers as H  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    import rootsolve_b200 as rs
    from rootsolve_b200 import parallel

    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n3 = int(os.environ.get("MGPU_N3", "48"))
    cases = [
        ("steady.3D %d^3 diff3d" % n3, lambda: rs.steady_3D(np.full(n3 ** 3, 10.0), func="diff3d", parms=H.P_DIFF3D, dimens=[n3] * 3,
                                                            positive=True, ipar=[1, n3, n3, n3, 0], **H.TOL)),
        ("steady.3D 20x18x%d rd3d (2 species)" % (2 * world + 3),
         lambda: rs.steady_3D(np.concatenate([np.full(20 * 18 * (2 * world + 3), 10.0), np.full(20 * 18 * (2 * world + 3), 1.0)]),
                              func="rd3d", parms=H.P_RD3D, nspec=2, dimens=[20, 18, 2 * world + 3], positive=True,
                              ipar=[2, 2 * world + 3, 18, 20, 0], **H.TOL)),
        ("steady.2D 101x67 Monod", lambda: rs.steady_2D(np.full(101 * 67, 10.0), func="diff2d", parms=H.P_DIFF2D_MONOD,
                                                        dimens=[101, 67], positive=True, ipar=[1, 67, 101, 1, 1], **H.TOL)),
        ("steady.1D 5 x 3001 rt1d", lambda: rs.steady_1D(np.full(5 * 3001, 1.0), nspec=5, method="stodes", func="rt1d",
                                                         parms=H.P_RT1D, positive=True, ipar=[5, 3001, 1, 1, 0], **H.TOL)),
    ]
    single = [f() for _, f in cases]          # before the communicator exists: plain single-GPU solves
    parallel.init_comm(dist)
    failures = 0
    for (name, f), ref in zip(cases, single):
        out = f()
        ok, worst = H.within_tol(out["y"], ref["y"])
        A, B = out.attributes, ref.attributes
        same = (A["dims"]["nnz"] == B["dims"]["nnz"] and A["dims"]["ngp"] == B["dims"]["ngp"] and A["steady"] and B["steady"])
        good = ok and same
        flag = torch.tensor([0 if good else 1], device="cuda")
        dist.all_reduce(flag)
        if rank == 0:
            print("%-40s ranks=%d  max err/(atol+rtol|y|)=%.3e  newton %d vs %d  krylov %d vs %d  %s" % (
                name, world, worst, len(A["precis"]), len(B["precis"]), A["krylov_iters"], B["krylov_iters"],
                "OK" if flag.item() == 0 else "FAIL"), flush=True)
        failures += int(flag.item() != 0)
    parallel.finalize_comm()
    dist.destroy_process_group()
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
