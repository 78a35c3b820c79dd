"""Fused tcgen05 GEMM -> reduce-scatter (csrc/gemm_rs.cu) against an fp32 matmul of the same bf16 operands.

The kernel was written while the pool's GPUs were closed, so these tests have never executed: they are opt-in
(``CML_EXPERIMENTAL=1``) until a first hardware run has confirmed them, to keep an unproven kernel out of the default GPU
suite.  Run each under a timeout - a protocol mistake in a spinning kernel shows up as a hang."""
import ctypes
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("CML_EXPERIMENTAL") != "1", reason="unverified kernel: set CML_EXPERIMENTAL=1")]


def _operands(p, B, K, seed):
    g = torch.Generator().manual_seed(seed)
    ld = (K + 7) // 8 * 8
    xt = torch.zeros(p, ld, dtype=torch.bfloat16, device="cuda")
    xt[:, :K] = torch.randn(p, K, generator=g).cuda().bfloat16()
    r = torch.zeros(B, ld, device="cuda")
    r[:, :K] = torch.randn(B, K, generator=g).cuda()
    return xt, r


@pytest.mark.parametrize("use_tma", [False, True])
@pytest.mark.parametrize("p,B,K", [(512, 256, 4096), (2000, 64, 10000), (4096, 512, 1000)])
def test_single_rank_matches_matmul(p, B, K, use_tma):
    from consensusml_b200.parallel import fused as F
    xt, r = _operands(p, B, K, 1)
    op = F.XtrGemmReduceScatter(p, B, "cuda", use_tma_reduce=use_tma)
    for rt in (r.bfloat16(), F.split_bf16(r)):
        for _ in range(2):                                   # twice: alternating buffers, increasing epochs
            op.launch(xt, rt, alpha=0.5)
            op.wait()
            torch.cuda.synchronize()
            got = op.result().clone()
        ref = 0.5 * F.xtr_gemm_rs_unfused(xt, rt, p, B, op.genes_per_rank)
        scale = max(1.0, float(ref.abs().max()))
        assert got.shape == ref.shape and (got - ref).abs().max().item() < 2e-3 * scale
    exact = 0.5 * (xt.float() @ r.t())                       # the hi/lo split recovers the fp32 residuals
    assert (got - exact).abs().max().item() < 5e-4 * max(1.0, float(exact.abs().max()))


def test_emulated_four_owners_on_one_gpu():
    """world = 4 with every 'peer' buffer on this GPU: each owner buffer must receive exactly its genes' rows."""
    from consensusml_b200 import _native
    from consensusml_b200.parallel import fused as F
    p, B, K, world = 1000, 128, 3000, 4
    xt, r = _operands(p, B, K, 2)
    rt = r.bfloat16()
    gpr = ((p + world - 1) // world + 31) // 32 * 32
    outs = [torch.zeros(gpr, B, device="cuda") for _ in range(world)]
    sigs = [torch.zeros(256, dtype=torch.int32, device="cuda") for _ in range(world)]
    done = torch.zeros(4, dtype=torch.int32, device="cuda")
    arr = lambda ts: (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    rc = _native.cuda().cml_xtr_gemm_reduce_scatter(p, B, B, xt.shape[1], _native.ptr(xt), xt.stride(0), _native.ptr(rt), rt.stride(0),
                                                    arr(outs), B, gpr, arr(sigs), 2, world, _native.ptr(done), 7, F.SIG_GEMM_RS, 1.0, 0,
                                                    _native.stream_ptr())
    _native.check(rc, "cml_xtr_gemm_reduce_scatter")
    torch.cuda.synchronize()
    ref = xt.float() @ rt.float().t()
    for d in range(world):
        lo, hi = min(p, d * gpr), min(p, (d + 1) * gpr)
        assert (outs[d][: hi - lo] - ref[lo:hi]).abs().max().item() < 2e-3 * float(ref.abs().max())
        assert int(sigs[d][F.SIG_GEMM_RS + 2]) == 7           # rank 2 signalled every destination
