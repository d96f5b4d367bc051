"""Multi-GPU parity (needs >= 2 CUDA devices; skipped on a 1-GPU box): the row-sharded evaluation
over NCCL equals the single-GPU evaluation.  Launches tests/sharded_gpu_worker.py under torchrun."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("peer", ["1", "0"])
def test_sharded_equals_single_gpu(peer):
    """peer=1: induced-field exchange inside the update kernel over CUDA-IPC peer memory (default);
    peer=0: NCCL all-reduce between the contraction and the update."""
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n < 4 else 4
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tests", "sharded_gpu_worker.py")], capture_output=True, text=True, timeout=900,
                       env=dict(os.environ, MPMC_B200_PEER=peer))
    assert f"PEER_ACTIVE={peer}" in r.stdout, r.stdout[-4000:] + r.stderr[-4000:]
    assert r.returncode == 0 and "SHARDED_OK" in r.stdout, r.stdout[-4000:] + r.stderr[-4000:]
