"""N > 1 on real GPUs: scripts/multi_gpu_check.py under torchrun, one rank per GPU, NCCL halo exchange and allreduce inside
libddm_b200.so; assembled results against the single-process oracle.  Skipped when fewer than two GPUs are visible
(the host-side sharding logic is covered on CPU by test_host.py and test_dist_gloo.py)."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_solve_matches_oracle(world):
    if _ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "scripts", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "FAIL" not in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_rank_local_setup_path_matches_oracle(world):
    """bench.py's default path: rank-local setup, device factorisation, device GenEO, joint Galerkin product (world = 1 runs on
    every GPU box, so the path is covered even where only one GPU is visible)."""
    if _ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "scripts", "multi_gpu_local_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "FAIL" not in r.stdout and "ok" in r.stdout
