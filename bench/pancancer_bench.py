#!/usr/bin/env python
"""BASELINE.json config 3 - pan-cancer multiclass run on a synthetic count matrix of the reference's pan-cancer shape
(60,488 genes x 474 samples, 4 classes; the real matrix is stripped from the reference tree): TMM / log2-CPM
normalisation, one-vs-rest limma-voom prefilter, the one-vs-rest ensemble (lasso with k-fold CV x exclusion replicates,
linear + RBF SVM, random forest, boosted trees) and the pan-class consensus ranking, on one GPU.  Wall clock per stage
(host timer around synchronised device work) and held-out accuracy; one JSON line."""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from consensusml_b200.runtime import clocks as CLK

CLK.start_global()          # SM clocks / throttle reasons are sampled while every measurement below runs

from consensusml_b200 import pipeline
from consensusml_b200.io.synthetic import nb_counts_multiclass
from consensusml_b200.runtime import config as C


def main():
    cpu = "--cpu" in sys.argv                      # dry run of the script logic without a GPU (use with --small)
    dev = torch.device("cpu" if cpu else "cuda")
    small = "--small" in sys.argv
    genes, sizes = (6000, (40, 40, 30, 30)) if small else (60488, (145, 160, 70, 99))
    t0 = time.perf_counter()
    counts, label = nb_counts_multiclass(genes, sizes, seed=11)
    gen_s = time.perf_counter() - t0
    rng = np.random.default_rng(2019)
    train = np.ones(label.size, bool)
    for c in np.unique(label):                                   # stratified 2/3 - 1/3 split (DEL:159-186)
        idx = rng.permutation(np.flatnonzero(label == c))
        train[idx[: len(idx) // 3]] = False
    out_dir = tempfile.mkdtemp(prefix="pancancer_")
    cfg = C.load(None, ["io.out_dir=" + out_dir, ("device=cpu" if cpu else "device=cuda"), "lasso.exclusion_reps=3", "lasso.nfolds=5", "rf.n_trees=[2000]",
                        'gbm.reps=[{"max_depth": 3, "nrounds": 50}]', "mlp.enabled=false"])
    c_dev = torch.from_numpy(counts).to(dev)
    ids = [f"ENSG{i:011d}.1" for i in range(genes)]
    sync = (lambda: None) if cpu else torch.cuda.synchronize
    sync(); t0 = time.perf_counter()
    co = pipeline.multiclass_cohort_from_counts(c_dev, ids, [f"s{i}" for i in range(label.size)], label, train, cfg)
    sync(); prep_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    s = pipeline.run_multiclass(cfg, co, class_names={0: "AML", 1: "NBL", 2: "WT", 3: "RT"})
    sync(); fit_s = time.perf_counter() - t0
    print(json.dumps({"clocks": CLK.snapshot(), "config": "pan-cancer multiclass (synthetic counts of the reference's shape)", "genes": genes, "samples": int(label.size),
                      "classes": s["classes"], "prefilter_genes": co.n_genes, "normalise_plus_4x_voom_s": round(prep_s, 3),
                      "one_vs_rest_ensemble_plus_consensus_s": round(fit_s, 3), "total_s": round(prep_s + fit_s, 3),
                      "heldout_accuracy": s.get("accuracy"), "confusion": s.get("confusion"), "top10": s.get("top", [])[:10],
                      "synthetic_generation_s_not_counted": round(gen_s, 2)}))


if __name__ == "__main__":
    main()
