"""The multi-GPU check of the fused compute + collective paths (tests/dist_fused_check.py) WITHOUT GPUs: one process per rank
under torchrun / gloo, every kernel on the host execution model, the symmetric buffers in POSIX shared memory that all ranks
map (parallel/symm.py), so the in-kernel peer stores, reductions and release / acquire signals really cross process
boundaries and the ranks really run concurrently.  Covers the Python orchestration around the kernels as well:
ShardTranspose, XtrReduceScatter (atomic and deterministic), the staged / direct all-gather Gram, HistAllReduce and the
HistEngine reduce-scatter landing, lasso_large (lasso, elastic net, batched lanes through the fused GEMM reduce-scatter),
GBMLarge's kernel level loop, the working-set SVM."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world,port", [(2, 29671), (3, 29672), (8, 29678)])
def test_fused_paths_between_emulated_ranks(world, port):
    """(The 3-rank run also switches on CML_SYMM_POISON: staging buffers are NaN until this epoch's copiers have flagged them.)"""
    if world > (os.cpu_count() or 1):
        pytest.skip("fewer cores than ranks")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "dist_fused_check.py"), "--emu"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1500, cwd=ROOT, env={**os.environ, "CML_DIST_TIMEOUT_S": "600", "CML_SYMM_POISON": "1" if world == 3 else "0"})
    assert r.returncode == 0 and "ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("EXPERIMENTAL verdict: OK") == world


def test_fused_paths_bench_runs_on_emulated_ranks():
    """bench/fused_paths.py::measure (every path, opt-in ones included) as two emulated ranks: the bench code itself is exercised."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29683", os.path.join(ROOT, "tests", "dist_bench_smoke.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1500, cwd=ROOT, env={**os.environ, "CML_DIST_TIMEOUT_S": "600"})
    assert r.returncode == 0 and "BENCH SMOKE OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_a_silent_peer_makes_the_wait_kernel_trap():
    """Failure detection: a rank that never signals turns into an error on the waiting rank after CML_WAIT_TIMEOUT_S, not a hang."""
    import time
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29713", os.path.join(ROOT, "tests", "dist_fault_check.py")]
    t0 = time.time()
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    out = r.stdout + r.stderr
    assert r.returncode != 0 and "__trap() reached" in out and "WAIT RETURNED" not in out, out[-3000:]
    assert "waiting for a peer that will never push" in out and time.time() - t0 < 120
