"""The whole pipeline on a GPU, from raw counts and from the bundled cohort."""
import json
import os

import numpy as np
import pytest
import torch

from consensusml_b200 import pipeline
from consensusml_b200.runtime import config as C

pytestmark = pytest.mark.gpu


def test_pipeline_reference_cohort_gpu(tmp_path):
    cfg = C.load(None, ["io.out_dir=" + str(tmp_path), "lasso.exclusion_reps=3", "rf.n_trees=[2000]", "mlp.enabled=true", "mlp.epochs=100"])
    s = pipeline.run(cfg)
    assert s["lasso"]["confusion"] == [[21, 2], [1, 20]] and abs(s["lasso"]["auc"] - 0.9835) < 0.01
    assert s["lasso"]["n_genes"][:3] == [13, 15, 24]                        # exclusion replicates 1-3 of the reference
    assert s["xgb1"]["n_genes"] == 6 and s["xgb2"]["n_genes"] == 7           # xgb_resultslist.rda importance tables
    assert 0.04 <= s["rf_2k"]["oob_error"] <= 0.15
    assert s["svm1"]["auc"] > 0.9 and s["mlp"]["auc"] > 0.85
    assert len(s["consensus"]["top"]) == 50
    recs = [json.loads(l) for l in open(tmp_path / "run.jsonl")]
    assert {r.get("group") for r in recs if r["event"] == "task"} >= {"lasso", "svm", "rf_2k", "gbm", "mlp"}
    assert any(r["event"] == "stage" and r.get("stage") == "write_outputs" for r in recs)


def test_counts_to_cohort_gpu():
    from consensusml_b200.io.synthetic import nb_counts
    c, g = nb_counts(20000, 96, seed=2, n_de=800)
    train = np.arange(96) % 3 != 0
    ids = [f"ENSG{i:011d}.1" for i in range(20000)]
    co_g = pipeline.cohort_from_counts(torch.from_numpy(c).cuda(), ids, [f"s{i}" for i in range(96)], g, train, C.Config())
    co_c = pipeline.cohort_from_counts(torch.from_numpy(c), ids, [f"s{i}" for i in range(96)], g, train, C.Config())
    a, b = set(co_g.gene_ids), set(co_c.gene_ids)
    assert len(a ^ b) <= max(3, 0.02 * len(b))
    common = [i for i in co_c.gene_ids if i in a]
    ia = [co_g.gene_ids.index(i) for i in common[:200]]
    ib = [co_c.gene_ids.index(i) for i in common[:200]]
    assert (co_g.assay[ia].cpu() - co_c.assay[ib]).abs().max() < 2e-3


@pytest.mark.gpu
def test_overlapped_stages_on_cuda_streams(tmp_path):
    """dist.task_streams: the stages of one rank on concurrent CUDA streams (host threads) - same lasso result, same genes on top."""
    from consensusml_b200 import pipeline
    from consensusml_b200.runtime import config as C
    ov = ["device=cuda:0", "lasso.exclusion_reps=3", "rf.n_trees=[500, 1000]", "checkpoint=false", "write_outputs=false",
          'gbm.reps=[{"max_depth": 2, "nrounds": 20}, {"max_depth": 3, "nrounds": 20}]']
    a = pipeline.run(C.load(None, ov + ["io.out_dir=" + str(tmp_path / "seq")]))
    b = pipeline.run(C.load(None, ov + ["io.out_dir=" + str(tmp_path / "par"), "dist.task_streams=4"]))
    assert "failed_tasks" not in a and "failed_tasks" not in b
    assert a["lasso"] == b["lasso"]
    assert len(set(a["consensus"]["top"]) & set(b["consensus"]["top"])) >= 45      # forests use float atomics: order of ties may move
