#!/usr/bin/env python
"""BASELINE.json config 5 - stability selection on the TARGET-AML training cohort (93 x 1984): 1000 bootstrap
replicates x elastic-net alpha grid {0.25, 0.5, 0.75, 1.0} x the reference's 100-lambda grid (400,000 lambda-point
fits in 4,000 path fits), plus an RBF-SVM (C, gamma) grid with 5-fold CV.  GPU wall clock (synchronised) and the CPU
arm's rate measured on a subset with the float64 C++ restatement of glmnet (one core), extrapolated; one JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from consensusml_b200.runtime import clocks as CLK

CLK.start_global()          # SM clocks / throttle reasons are sampled while every measurement below runs

from consensusml_b200 import datasets
from consensusml_b200.models import lasso as L
from consensusml_b200.models import svm as S
from consensusml_b200.ops import lasso as K
from consensusml_b200.ops.replicate import replicate_weights


def main():
    cpu = "--cpu" in sys.argv                      # dry run of the script logic without a GPU (use with --small)
    n_boot = 8 if "--small" in sys.argv else 1000
    dev = torch.device("cpu" if cpu else "cuda")
    sync = (lambda: None) if cpu else torch.cuda.synchronize
    co = datasets.load_target_aml()
    tr, _ = co.split()
    x, y = tr.design(torch.float64 if cpu else torch.float32).to(dev), tr.labels().to(dev).to(torch.float64 if cpu else torch.float32)
    lam = K.default_lambda_grid()
    alphas = (0.25, 0.5, 0.75, 1.0)
    L.stability_selection(x, y, lam, n_boot=min(64, n_boot), alphas=(1.0,))         # warm-up
    sync(); t0 = time.perf_counter()
    out = L.stability_selection(x, y, lam, n_boot=n_boot, alphas=alphas, seed=1)
    sync(); stab_s = time.perf_counter() - t0
    top = torch.argsort(-out["max_freq"])[:10].cpu().tolist()
    costs, gammas = [0.25, 1.0, 4.0, 16.0], [0.25 / x.shape[1], 1.0 / x.shape[1], 4.0 / x.shape[1]]
    S.grid_search(x.float(), y.float(), costs[:1], gammas[:1])
    sync(); t0 = time.perf_counter()
    gs = S.grid_search(x.float(), y.float(), costs, gammas, nfolds=5)
    sync(); svm_s = time.perf_counter() - t0
    # CPU arm: 8 bootstrap path fits per alpha on one core
    xc, yc = x.double().cpu(), y.double().cpu()
    cpu_rate = {}
    for a in alphas:
        w = replicate_weights(8, xc.shape[0], seed=3).double()
        t0 = time.perf_counter()
        K.lognet_path(xc, yc, w, lam, alpha=a, want_eta=False)
        cpu_rate[a] = 8 / (time.perf_counter() - t0)
    cpu_est = sum(n_boot / r for r in cpu_rate.values())
    print(json.dumps({"clocks": CLK.snapshot(), "config": "stability selection: 1000 bootstraps x alpha grid + RBF-SVM grid (TARGET-AML 93 x 1984)",
                      "path_fits": 4 * n_boot, "gpu_stability_s": round(stab_s, 3), "gpu_path_fits_per_s": round(4 * n_boot / stab_s, 1),
                      "gpu_svm_grid_s": round(svm_s, 3), "svm_grid_points": len(costs) * len(gammas) * 5,
                      "svm_best": [gs["best_cost"], gs["best_gamma"]], "cpu_one_core_path_fits_per_s": {str(k): round(v, 2) for k, v in cpu_rate.items()},
                      "cpu_one_core_stability_s_extrapolated": round(cpu_est, 1), "speedup_vs_one_core": round(cpu_est / stab_s, 1),
                      "top10_genes_by_max_selection_frequency": [co.gene_ids[i] for i in top],
                      "max_freq_top10": [round(float(out["max_freq"][i]), 3) for i in top]}))


if __name__ == "__main__":
    main()
