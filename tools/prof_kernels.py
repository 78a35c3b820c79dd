"""Launch each hot kernel once at benchmark shape (after a warm-up launch) so ncu can capture it."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from consensusml_b200.ops import gemm as G
from consensusml_b200.parallel import fused as F
which = sys.argv[1] if len(sys.argv) > 1 else "all"
dev = torch.device("cuda")
if which in ("gemm", "all"):
    M = N = 8192; K = 8192
    a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
    na = torch.ones(M, device=dev); nb = torch.ones(N, device=dev); c = torch.empty(M, N, device=dev)
    for _ in range(2):
        G.gemm_tn(a, [b], N, epilogue=G.EPI_RBF, alpha=1e-4, norm_a=na, norm_b=[nb], out=c)
    torch.cuda.synchronize(); print("gemm done")
if which in ("xtr", "all"):
    n, p = 125000, 20000
    xb = torch.empty(n, p, dtype=torch.bfloat16, device=dev).normal_(); R = torch.randn(n, 1, device=dev)
    rs = F.XtrReduceScatter(p, 1, dev)
    for _ in range(2):
        rs.launch(xb, R); rs.wait(); torch.cuda.synchronize(); rs.grad.local.zero_()
    print("xtr done")
if which in ("hist", "all"):
    n, p = 125000, 20000
    bins = torch.randint(0, 256, (p, n), dtype=torch.uint8, device=dev)
    node = torch.randint(0, 4, (n,), device=dev).to(torch.int16); gh = torch.randn(n, 2, device=dev)
    H = F.HistAllReduce(p, 256, 4, dev)
    for _ in range(2):
        H.build(bins, n, node, gh, 4, use_multicast=False); H.clear(4)
    print("hist done")
