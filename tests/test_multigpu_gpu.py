"""Two ranks over NCCL on a box with at least two GPUs (skipped elsewhere; the one-GPU emulation of the same logic is
tests/test_fused_step_gpu.py, the CPU/gloo one tests/test_fused_shard_cpu.py): profiles/multigpu_check.py under torchrun --
sharded direct and Barnes-Hut evaluations, and fused sharded trajectories, against the one-GPU results bit for bit."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_ranks_equal_one_gpu_bit_for_bit():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cp = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                         "--master-addr", "127.0.0.1", "--master-port", str(port),
                         os.path.join(ROOT, "profiles", "multigpu_check.py")], capture_output=True, text=True, timeout=600)
    assert cp.returncode == 0, cp.stdout[-2000:] + cp.stderr[-2000:]
    assert "MULTIGPU_CHECK PASS world 2" in cp.stdout, cp.stdout[-2000:]
