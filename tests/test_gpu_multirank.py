"""Two ranks, two GPUs, NCCL tile broadcasts: slate::gemm and slate::potrf on 1x2 and 2x1
grids (device-origin and ScaLAPACK-layout host buffers) against the CPU oracle.  Spawns
scripts/mp_parity.py under torchrun; skipped on a box with a single GPU (the same script
is what the 8-GPU runs use)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_gpus_nccl_parity():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "scripts", "mp_parity.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0 and "MP_PARITY PASS" in out.stdout, out.stdout[-3000:] + out.stderr[-2000:]
