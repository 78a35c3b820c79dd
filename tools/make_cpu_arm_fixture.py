#!/usr/bin/env python
"""Run the default consensus pipeline on the CPU (float64 host solvers / NumPy tree restatement, one process) and store
what bench.py compares a GPU run against: per-model held-out AUCs, the lasso gene set and the consensus ranking.
Written to consensusml_b200/datasets/target_aml/cpu_arm_summary.json together with the wall-clock it took here."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from consensusml_b200 import datasets, pipeline
from consensusml_b200.runtime import config as C

torch.set_num_threads(1)                      # the reference ran single-threaded (n_jobs=1)
cfg = C.load(None, ["device=cpu", "resume=false", "checkpoint=false", "write_outputs=false", "io.out_dir=/tmp/cml_cpu_arm"] + sys.argv[1:])
t0 = time.perf_counter()
det = pipeline.run(cfg, return_details=True)
wall = time.perf_counter() - t0
s = det["summary"]
out = {"config_hash": cfg.hash(["lasso", "svm", "rf", "gbm", "mlp", "split", "consensus"]), "wall_s_one_core": wall,
       "consensus_ranking": det["rank"].index.tolist()[:200], "lasso_genes_rep1": s["lasso"]["genes_rep1"],
       "lasso": {k: s["lasso"][k] for k in ("auc", "confusion", "lambda_min", "n_genes", "test_errors")},
       "auc": {k: v["auc"] for k, v in s.items() if isinstance(v, dict) and "auc" in v},
       "group_seconds": s["schedule"]["group_seconds"]}
path = os.path.join(ROOT, "consensusml_b200", "datasets", "target_aml", "cpu_arm_summary.json")
json.dump(out, open(path, "w"), indent=1)
print(json.dumps({k: out[k] for k in ("wall_s_one_core", "auc", "group_seconds")}))
