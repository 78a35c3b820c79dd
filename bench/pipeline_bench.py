#!/usr/bin/env python
"""Wall-clock of the TARGET-AML risk-group ensemble (the cml_targetaml_seanalysis.Rmd run: lasso x15 exclusion
replicates with LOOCV, SVM x4, random forest 2k/5k/10k trees with proximity, xgboost x5) on one B200 versus a CPU
arm built from compiled CPU libraries on the same data: the float64 C++ restatement of glmnet (libcml_host.so, validated
to 1e-15 against the reference's stored glmnet objects), scikit-learn's libsvm SVC, RandomForestClassifier and
GradientBoostingClassifier (xgboost itself is not installed; the substitution is recorded in the output).
Prints one JSON line."""
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
from consensusml_b200.models import trees as T
from consensusml_b200.ops import metrics as M


def gpu_arm(co, dev):
    tr, te = co.split()
    xtr, ytr, xte, yte = tr.design(torch.float32).to(dev), tr.labels().to(dev), te.design(torch.float32).to(dev), te.labels().to(dev)
    out = {}
    torch.cuda.synchronize(); t0 = time.perf_counter()
    reps = L.exclusion_reps(co, nreps=15, device=dev)
    torch.cuda.synchronize(); out["lasso_15reps_loocv_s"] = time.perf_counter() - t0
    out["lasso_nnz"] = [len(r.nonzero_coef) for r in reps]
    out["lasso_test_err"] = [round(r.test_error, 3) for r in reps]
    out["lasso_auc_rep1"] = reps[0].test_auc
    t0 = time.perf_counter()
    aucs = []
    for kern in ("linear", "radial"):
        for wf in (None, 0.5):
            aucs.append(S.run_svm(xtr, ytr, xte, yte, kern, weightfilt=wf).auc)
    torch.cuda.synchronize(); out["svm_x4_s"] = time.perf_counter() - t0
    out["svm_auc"] = aucs
    t0 = time.perf_counter()
    rf_err = {}
    for nt in (2000, 5000, 10000):
        rf = T.RandomForest(n_trees=nt, proximity=True).fit(xtr, ytr)
        rf_err[nt] = (rf.oob_error_, float((rf.predict(xte) != yte.long()).float().mean()))
    torch.cuda.synchronize(); out["rf_2k5k10k_s"] = time.perf_counter() - t0
    out["rf_oob_test_err"] = rf_err
    t0 = time.perf_counter()
    xg = []
    for depth, rounds in ((2, 2), (50, 2), (50, 50), (100, 50), (100, 100)):
        gb = T.GradientBoosting(n_estimators=rounds, learning_rate=1.0, max_depth=depth).fit(xtr, ytr)
        xg.append(int((gb.importance_["gain"] > 0).sum()))
    torch.cuda.synchronize(); out["xgb_x5_s"] = time.perf_counter() - t0
    out["xgb_genes"] = xg
    out["total_s"] = out["lasso_15reps_loocv_s"] + out["svm_x4_s"] + out["rf_2k5k10k_s"] + out["xgb_x5_s"]
    return out


def cpu_arm(co, lasso_reps=15):
    from sklearn.ensemble import GradientBoostingClassifier, RandomForestClassifier
    from sklearn.svm import SVC
    tr, te = co.split()
    xtr, ytr = tr.design(torch.float64).numpy(), tr.labels().numpy()
    out = {}
    t0 = time.perf_counter()
    L.exclusion_reps(co, nreps=lasso_reps, device=None)                # C++ float64 glmnet restatement, one core
    out["lasso_15reps_loocv_s"] = (time.perf_counter() - t0) * 15 / lasso_reps
    t0 = time.perf_counter()
    xs = (xtr - xtr.mean(0)) / xtr.std(0, ddof=1)
    for kern in ("linear", "rbf"):
        for cols in (slice(None), slice(0, 992)):
            SVC(kernel=kern, gamma=1.0 / xs.shape[1]).fit(xs[:, cols], ytr)
    out["svm_x4_s"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    for nt in (2000, 5000, 10000):
        RandomForestClassifier(n_estimators=nt, max_features=44, n_jobs=1, oob_score=True, random_state=20).fit(xtr, ytr)
    out["rf_2k5k10k_s"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    for depth, rounds in ((2, 2), (50, 2), (50, 50), (100, 50), (100, 100)):
        GradientBoostingClassifier(n_estimators=rounds, learning_rate=1.0, max_depth=depth, random_state=0).fit(xtr, ytr)
    out["xgb_x5_s"] = time.perf_counter() - t0
    out["total_s"] = sum(v for k, v in out.items() if k.endswith("_s"))
    out["note"] = "single core (the reference ran n_jobs=1 / nthread=2); sklearn GradientBoosting stands in for xgboost; proximity not computed by sklearn RF"
    return out


if __name__ == "__main__":
    co = datasets.load_target_aml()
    dev = torch.device("cuda")
    gpu_arm(co, dev)                                   # warm-up: library load, kernel images, allocator
    g = gpu_arm(co, dev)
    c = cpu_arm(co, lasso_reps=int(os.environ.get("CML_CPU_LASSO_REPS", 3)))
    print(json.dumps({"clocks": CLK.snapshot(), "bench": "target_aml_ensemble", "gpu": g, "cpu": c, "speedup_total": c["total_s"] / g["total_s"],
                      "speedup": {k: c[k] / g[k] for k in c if k.endswith("_s") and k in g}}))
