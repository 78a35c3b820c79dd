"""Fused compute+collective paths: single-rank on one GPU (all peers local), 2 ranks when two GPUs are visible,
and the sharding logic under gloo on CPU."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = os.path.join(ROOT, "tests", "dist_fused_check.py")


def _torchrun(nproc, extra, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), SCRIPT] + extra
    return subprocess.run(cmd, capture_output=True, text=True, timeout=900)


def test_sharded_lasso_gloo_two_ranks_matches_single_process():
    r = _torchrun(2, ["--cpu"], 29541)
    assert r.returncode == 0 and "ALL OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.gpu
def test_fused_paths_single_rank():
    r = subprocess.run([sys.executable, SCRIPT], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.gpu
def test_fused_paths_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    r = _torchrun(2, [], 29543)
    assert r.returncode == 0 and "ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
