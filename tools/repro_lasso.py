import sys, os, subprocess, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
def one(B, eta, coef, freq, rw, seed):
    import torch, numpy as np, bench
    from consensusml_b200.ops import lasso as K
    from consensusml_b200.ops.replicate import replicate_weights
    x, y = bench.synth_cohort(seed)
    x = torch.from_numpy(x).cuda(); y = torch.from_numpy(y).cuda()
    lam = K.default_lambda_grid()
    packed = K.pack_design(x)
    if rw:
        w = replicate_weights(B, 93, seed=seed, device="cuda", n_pad=96)
    else:
        w = torch.ones(B, 96, device="cuda"); w[:, 93:] = 0
    torch.cuda.synchronize()
    fq = torch.zeros(100, 1984, device="cuda") if freq else None
    pb = K.lognet_path(x, y, w[:, :93], lam, want_eta=eta, want_coef=coef, freq_out=fq, packed=packed, padded_weights=w)
    torch.cuda.synchronize()
    print("ok", B, eta, coef, freq, rw, seed, "jerr", pb.jerr.unique().tolist(), "npasses max", int(pb.npasses.max()), flush=True)
if __name__ == "__main__":
    if len(sys.argv) > 1:
        a = sys.argv[1:]
        one(int(a[0]), a[1] == "1", a[2] == "1", a[3] == "1", a[4] == "1", int(a[5]))
    else:
        for cfg in [(64,1,1,0,0,0),(64,0,0,1,1,0),(1024,1,1,0,0,0),(1024,0,0,1,0,0),(1024,0,0,1,1,0),(1024,0,1,0,1,0),(4096,0,0,1,1,0),(4096,1,1,0,0,0),(4096,0,0,0,0,0),(4096,0,0,0,1,3)]:
            r = subprocess.run([sys.executable, __file__] + [str(c) for c in cfg], capture_output=True, text=True)
            print(cfg, "rc", r.returncode, r.stdout.strip()[-200:], r.stderr.strip()[-300:] if r.returncode else "", flush=True)
