#!/usr/bin/env python
"""tcgen05 GEMM throughput vs cuBLAS (torch.matmul bf16) on square and framework-shaped problems; CUDA events,
L2 flushed between repetitions.  One JSON line per shape."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from consensusml_b200.runtime import clocks as CLK

CLK.start_global()          # SM clocks / throttle reasons are sampled while every measurement below runs

from consensusml_b200 import _native
from consensusml_b200.ops import gemm as G

lib = _native.cuda()
fb = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
flush = lambda: lib.cml_l2_flush(_native.ptr(fb), fb.numel(), _native.stream_ptr())


def timed(fn, reps=7):
    fn(); torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return sorted(ms)[len(ms) // 2]


for (M, N, K) in [(4096, 4096, 4096), (8192, 8192, 8192), (16384, 16384, 2048), (8192, 8192, 20000), (1984, 1984, 144), (2048, 65536, 1024)]:
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    c = torch.empty(M, (N + 3) // 4 * 4, device="cuda")
    na = torch.ones(M, device="cuda"); nb = torch.ones(N, device="cuda")
    G.set_variant(1)
    one_cta = timed(lambda: G.gemm_tn(a, [b], N, out=c))
    G.set_variant(2)
    pair = timed(lambda: G.gemm_tn(a, [b], N, out=c))
    G.set_variant(0)
    ours = timed(lambda: G.gemm_tn(a, [b], N, out=c))
    rbf = timed(lambda: G.gemm_tn(a, [b], N, epilogue=G.EPI_RBF, alpha=1e-4, norm_a=na, norm_b=[nb], out=c))
    cublas = timed(lambda: torch.matmul(a, b.t()))
    fl = 2.0 * M * N * K
    print(json.dumps({"clocks": CLK.snapshot(), "M": M, "N": N, "K": K, "ours_ms": ours, "ours_tflops": fl / ours / 1e9, "ours_rbf_tflops": fl / rbf / 1e9,
                      "one_cta_tflops": fl / one_cta / 1e9, "cta_pair_tflops": fl / pair / 1e9,
                      "cublas_bf16_out_ms": cublas, "cublas_tflops": fl / cublas / 1e9}), flush=True)
