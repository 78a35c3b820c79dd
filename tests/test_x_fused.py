"""Fused compute+collective paths: single-rank on one GPU (all peers local), 2 / 4 / 8 ranks when that many GPUs are
visible, and the sharding logic under gloo on CPU.  (Collected late on purpose - see tests/test_y_kernels_gpu.py.)"""
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
@pytest.mark.parametrize("nproc", [2, 4, 8])
def test_fused_paths_multi_rank(nproc):
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    r = _torchrun(nproc, [], 29543 + nproc)
    assert r.returncode == 0 and "ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.gpu
def test_pipeline_two_ranks_nccl(tmp_path):
    """pipeline.run spread over two GPUs reproduces the single-GPU standouttable (tests/dist_pipeline_check.py --cuda)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29561", os.path.join(ROOT, "tests", "dist_pipeline_check.py"), str(tmp_path), "--cuda"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "PIPELINE DIST OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
