"""Slab-sharded solve on 2 GPUs against the single-GPU solve (skipped on a 1-GPU box)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def test_two_gpu_sharded_solve_matches_single_gpu(built):
    if built.rs_device_count() < 2:
        pytest.skip("needs 2 GPUs (run: gpurun --gpus 2)")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(here, "mgpu_parity.py")]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    print(p.stdout[-3000:])
    assert p.returncode == 0, p.stdout[-3000:]
