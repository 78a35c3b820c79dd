#!/usr/bin/env python
"""Per-stage device times of the non-GEMM kernel families on reference-shaped inputs (CUDA events, median of reps after a
warm-up): TMM / log-CPM normalisation and limma-voom DE on 60,488 x 474 synthetic counts (the pan-cancer shape), and on the
TARGET-AML cohort (93 x 1984): random forest 10k trees + proximity, xgboost-style boosting, the batched SMO grid, Spearman
correlation 1984 x 1984.  One JSON line per stage, with the bytes that must move and the fraction of the HBM roofline that
implies.  ``--cpu --small`` is a dry run of the script logic."""
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


def main():
    cpu, small = "--cpu" in sys.argv, "--small" in sys.argv
    dev = torch.device("cpu" if cpu else "cuda")
    reps = 1 if cpu else 5
    try:
        hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        hbm = 6650.0

    def timed(fn):
        fn()
        ms = []
        for _ in range(reps):
            if cpu:
                t0 = time.perf_counter(); fn(); ms.append((time.perf_counter() - t0) * 1e3)
            else:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize()
                ms.append(a.elapsed_time(b))
        return sorted(ms)[len(ms) // 2]

    def emit(stage, ms, min_bytes=None, **kw):
        rec = {"stage": stage, "ms": round(ms, 3), **kw}
        if min_bytes:
            rec["min_bytes"] = int(min_bytes)
            rec["frac_of_hbm_roofline"] = round(min_bytes / (hbm * 1e9) / (ms / 1e3), 4)
        rec["clocks"] = CLK.snapshot()
        print(json.dumps(rec), flush=True)

    from consensusml_b200 import datasets
    from consensusml_b200.io.synthetic import nb_counts
    from consensusml_b200.models import svm as S
    from consensusml_b200.models import trees as T
    from consensusml_b200.stats import correlation as C
    from consensusml_b200.stats import de as DE
    from consensusml_b200.stats import normalize as N

    G, Sn = (4000, 40) if small else (60488, 474)
    counts_np, group = nb_counts(G, Sn, seed=3, n_de=G // 30)
    counts = torch.from_numpy(counts_np).to(dev)
    ms = timed(lambda: N.tmm_normalize(counts, 1.0, max(2, int(0.05 * Sn))))
    emit("tmm_normalize (lib sizes, CPM filter, TMM factors, log2-CPM)", ms, min_bytes=G * Sn * 4 * 3, genes=G, samples=Sn)
    ms = timed(lambda: DE.voom_de(counts, group))
    emit("voom_de (filter, TMM, voom weights, WLS, eBayes, BH)", ms, min_bytes=G * Sn * 4 * 4, genes=G, samples=Sn)

    co = datasets.load_target_aml()
    tr, te = co.split()
    xtr, ytr = tr.design(torch.float32).to(dev), tr.labels().to(dev)
    nt = 200 if small else 10000
    ms = timed(lambda: T.RandomForest(n_trees=nt, proximity=True).fit(xtr, ytr))
    emit(f"random forest {nt} trees + OOB + importance + proximity (93 x 1984)", ms, trees_per_s=round(nt / ms * 1e3))
    nr = 5 if small else 100
    ms = timed(lambda: T.GradientBoosting(n_estimators=nr, learning_rate=0.3, max_depth=6).fit(xtr, ytr))
    emit(f"xgboost-style boosting {nr} rounds depth 6 (93 x 1984)", ms, rounds_per_s=round(nr / ms * 1e3, 1))
    costs, gammas = ([1.0], [1.0 / 1984]) if small else ([0.25, 1.0, 4.0, 16.0], [0.25 / 1984, 1.0 / 1984, 4.0 / 1984])
    ms = timed(lambda: S.grid_search(xtr, ytr, costs, gammas, nfolds=5))
    emit(f"RBF-SVM grid {len(costs)} x {len(gammas)} x 5 folds (batched SMO)", ms, fits_per_s=round(len(costs) * len(gammas) * 5 / ms * 1e3, 1))
    x = co.assay.float().to(dev)
    x = x[:300] if small else x
    ms = timed(lambda: C.correlation(x))
    emit(f"Spearman correlation {x.shape[0]} x {x.shape[0]} over {x.shape[1]} samples (rank + standardise + GEMM)", ms,
         min_bytes=x.numel() * 4 + x.shape[0] ** 2 * 4)


if __name__ == "__main__":
    main()
