"""Multi-process, multi-GPU count: needs at least two CUDA devices (gpurun --gpus 2);
skipped on a one-GPU box, where test_gpu_parity.py covers the same kernels with shards in one process."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _devices():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("transport", ["peer", "nccl"])
def test_two_ranks_route_and_count(transport):
    if _devices() < 2:
        pytest.skip("needs two GPUs")
    n = min(_devices(), 4)
    n = 4 if n >= 4 else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", "29533" if transport == "peer" else "29534", os.path.join(ROOT, "tests", "mp_route_worker.py"), transport]
    p = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    out = p.stdout.decode()
    assert p.returncode == 0 and "mp_route_worker ok" in out, out[-3000:]
