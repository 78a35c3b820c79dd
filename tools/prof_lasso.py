"""One warm-up + one measured launch of the batched path kernel on a bench-shaped cohort (for ncu)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
from consensusml_b200.ops import lasso as K
from consensusml_b200.ops.replicate import replicate_weights
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4736
kopt = int(sys.argv[2]) if len(sys.argv) > 2 else 0
x, y = bench.synth_cohort(0)
x = torch.from_numpy(x).cuda(); y = torch.from_numpy(y).cuda()
lam = K.default_lambda_grid()
packed = K.pack_design(x)
w = replicate_weights(B, 93, seed=1, device="cuda", n_pad=96)
for it in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    pb = K.lognet_path(x, y, w[:, :93], lam, want_eta=False, want_coef=False, packed=packed, padded_weights=w, modified_newton=bool(kopt))
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"B={B} kopt={kopt} iter {it}: {dt*1e3:.1f} ms  -> {B/dt:.0f} fits/s; npasses mean {pb.npasses.float().mean():.0f}", flush=True)
