"""Multi-GPU parity of the N-sharded (column-parallel) path, SURVEY.md 8(e): G device shards + NCCL all-gather must
reproduce the single-GPU E bit for bit.  One process per GPU through torchrun (tests/shard_worker.py); needs >= 2 GPUs
(run with `gpurun --gpus 2|8 -- python -m pytest tests/test_gpu_shard.py -m gpu`), skipped on a 1-GPU box."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_gemm_plus_gather_equals_single_gpu(world):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "shard_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    sys.stdout.write(r.stdout[-4000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "gathered==single True" in r.stdout
