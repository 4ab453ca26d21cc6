"""The direct exchange with REAL ranks: one process per GPU under torchrun, receive windows
mapped with CUDA IPC, merged output against the oracle. Needs two GPUs on the box (skipped
otherwise; tests/test_gpu_exchange.py covers the same kernels with emulated ranks on one GPU)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    import torch

    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4])
def test_real_ranks_over_cuda_ipc(world):
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + world),
           os.path.join(ROOT, "tests", "_mp_exchange_worker.py")]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert p.stdout.count("MP_EXCHANGE OK mode direct") == 2, p.stdout[-2000:]   # both workloads
    print(p.stdout[-600:])
