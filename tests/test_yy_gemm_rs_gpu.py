"""Fused tcgen05 GEMM -> reduce-scatter (csrc/gemm_rs.cu) against an fp32 matmul of the same bf16 operands, on the GPU.

The kernel has run on the host model of the asynchronous units only (tests/test_emu_gemm.py: same source); its first hardware execution happens here (the model ran smaller shapes of every mode: both landings, hi / lo split, split-K, several owners).  Every case therefore runs in a child process under a
timeout: a protocol mistake in a spinning kernel would be a hang, and a hang must cost one failed test, not the session.
Collected next to last for the same reason (only the allocation-free entry point's test comes after it)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_COMMON = r"""
import ctypes, sys, torch
sys.path.insert(0, %r)
from consensusml_b200 import _native
from consensusml_b200.parallel import fused as F

def operands(p, B, K, seed):
    g = torch.Generator().manual_seed(seed)
    ld = (K + 7) // 8 * 8
    xt = torch.zeros(p, ld, dtype=torch.bfloat16, device="cuda")
    xt[:, :K] = torch.randn(p, K, generator=g).cuda().bfloat16()
    r = torch.zeros(B, ld, device="cuda")
    r[:, :K] = torch.randn(B, K, generator=g).cuda()
    return xt, r
""" % ROOT

_SINGLE = _COMMON + r"""
p, B, K, use_tma = %d, %d, %d, %d
xt, r = operands(p, B, K, 1)
op = F.XtrGemmReduceScatter(p, B, "cuda", use_tma_reduce=bool(use_tma))
for rt in (r.bfloat16(), F.split_bf16(r)):
    for _ in range(2):                                   # twice: alternating buffers, increasing epochs
        op.launch(xt, rt, alpha=0.5)
        op.wait()
        torch.cuda.synchronize()
        got = op.result().clone()
    ref = 0.5 * F.xtr_gemm_rs_unfused(xt, rt, p, B, op.genes_per_rank)
    scale = max(1.0, float(ref.abs().max()))
    assert got.shape == ref.shape and (got - ref).abs().max().item() < 2e-3 * scale, (got - ref).abs().max().item()
exact = 0.5 * (xt.float() @ r.t())                       # the hi/lo split recovers the fp32 residuals
assert (got - exact).abs().max().item() < 5e-4 * max(1.0, float(exact.abs().max()))
print("CHILD OK")
"""

_OWNERS = _COMMON + r"""
p, B, K, world = 1000, 128, 3000, 4
xt, r = operands(p, B, K, 2)
rt = r.bfloat16()
gpr = ((p + world - 1) // world + 31) // 32 * 32
outs = [torch.zeros(gpr, B, device="cuda") for _ in range(world)]
sigs = [torch.zeros(256, dtype=torch.int32, device="cuda") for _ in range(world)]
done = torch.zeros(4, dtype=torch.int32, device="cuda")
arr = lambda ts: (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
rc = _native.cuda().cml_xtr_gemm_reduce_scatter(p, B, B, xt.shape[1], _native.ptr(xt), xt.stride(0), _native.ptr(rt), rt.stride(0),
                                                arr(outs), B, gpr, arr(sigs), 2, world, _native.ptr(done), 7, F.SIG_GEMM_RS, 1.0, %d,
                                                _native.stream_ptr())
_native.check(rc, "cml_xtr_gemm_reduce_scatter")
torch.cuda.synchronize()
ref = xt.float() @ rt.float().t()
for d in range(world):
    lo, hi = min(p, d * gpr), min(p, (d + 1) * gpr)
    assert (outs[d][: hi - lo] - ref[lo:hi]).abs().max().item() < 2e-3 * float(ref.abs().max())
    assert int(sigs[d][F.SIG_GEMM_RS + 2]) == 7           # rank 2 signalled every destination
print("CHILD OK")
"""


_BATCHED = _COMMON + r"""
import numpy as np
from consensusml_b200.models import lasso_large as LL
g = torch.Generator().manual_seed(5)
n, p, B = 3000, 2500, 96
x = torch.randn(n, p, generator=g).bfloat16().float()
bt = torch.zeros(p); bt[:10] = torch.randn(10, generator=g)
y = (torch.rand(n, generator=g) < torch.sigmoid(x @ bt)).float()
lmax = LL.lambda_max(x, y, n)
lams = np.geomspace(lmax * 0.8, lmax * 0.05, B)
W = torch.poisson(torch.ones(B, n), generator=g)
ref = LL.fit_batched(x, y, lams, weights=W, backend="unfused", max_iter=50, tol=0.0)
got = LL.fit_batched(x.cuda(), y.cuda(), lams, weights=W.cuda(), backend="fused", max_iter=50, tol=0.0, step=ref.step, use_tma_reduce=bool(%d))
err = (got.beta.cpu() - ref.beta).abs().max().item()
assert err < 5e-4, err
assert (got.objective - ref.objective).abs().max().item() < 1e-4
print("CHILD OK")
"""


def _child(code: str, timeout: int = 240):
    try:
        res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    except subprocess.TimeoutExpired as exc:                 # subprocess.run has killed exactly the child it started
        pytest.fail(f"the kernel did not finish within {timeout} s (hang): {(exc.stderr or b'')[-2000:]!r}")
    assert res.returncode == 0 and "CHILD OK" in res.stdout, (res.stdout[-2000:], res.stderr[-4000:])


@pytest.mark.parametrize("use_tma", [0, 1])
@pytest.mark.parametrize("p,B,K", [(512, 256, 4096), (2000, 64, 10000), (4096, 512, 1000)])
def test_single_rank_matches_matmul(p, B, K, use_tma):
    _child(_SINGLE % (p, B, K, use_tma))


@pytest.mark.parametrize("use_tma", [0, 1])
def test_four_owners_on_one_gpu(use_tma):
    """world = 4 with every 'peer' buffer on this GPU: each owner buffer must receive exactly its genes' rows."""
    _child(_OWNERS % use_tma)


@pytest.mark.parametrize("use_tma", [0, 1])
def test_batched_sharded_lasso_on_the_kernels(use_tma):
    """BatchedLassoLarge (tcgen05 eta product, residual split, fused GEMM -> reduce-scatter, proximal step + all-gather)
    against the torch iteration of the same 96 lanes."""
    _child(_BATCHED % use_tma, timeout=400)


@pytest.mark.parametrize("nproc", [1, 2, 8])
def test_opt_in_fused_paths_on_gpus(nproc):
    """The section of tests/dist_fused_check.py that was written without hardware (fused GEMM -> reduce-scatter across ranks, the
    deterministic reduce-scatter, grouped staged order, sibling subtraction on the fused histogram path, elastic net, the
    batched sharded lasso) - passed with 2 - 8 emulated ranks; here on real GPUs, in its own process group and its own test."""
    import torch
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    script = os.path.join(ROOT, "tests", "dist_fused_check.py")
    cmd = ([sys.executable, script, "--only-experimental"] if nproc == 1 else
           [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
            "--master-port", str(29570 + nproc), script, "--only-experimental"])
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    except subprocess.TimeoutExpired as exc:
        pytest.fail(f"did not finish within 600 s: {(exc.stdout or b'')[-2000:]!r}")
    assert r.returncode == 0 and "ALL OK" in r.stdout and "EXPERIMENTAL verdict: OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
