import json
import os

import numpy as np
import pytest
import torch

from consensusml_b200 import pipeline
from consensusml_b200.runtime import config as C
from consensusml_b200.runtime.log import NumericalFailure, RunLog, check_finite, describe_jerr
from consensusml_b200.runtime.store import ArtifactStore


def test_config_defaults_overrides_and_hash(tmp_path):
    cfg = C.load(None, ["lasso.exclusion_reps=2", "rf.n_trees=[50]", "svm.enabled=false", "io.out_dir=" + str(tmp_path)])
    assert cfg.lasso.exclusion_reps == 2 and cfg.rf.n_trees == [50] and cfg.svm.enabled is False
    assert cfg.split.seed == 2019 and cfg.rf.seed == 20 and cfg.svm.seed == 50          # the reference's seeds
    h = cfg.hash()
    cfg.lasso.alpha = 0.5
    assert cfg.hash() != h and cfg.hash(["rf"]) == C.load(None, ["rf.n_trees=[50]"]).hash(["rf"])
    with pytest.raises(KeyError):
        C.load(None, ["lasso.nope=1"])
    p = tmp_path / "c.yaml"
    p.write_text("lasso:\n  alpha: 0.25\nconsensus:\n  method: borda\n")
    c2 = C.load(str(p))
    assert c2.lasso.alpha == 0.25 and c2.consensus.method == "borda"


def test_store_resume_and_log(tmp_path):
    log = RunLog(str(tmp_path / "run.jsonl"))
    store = ArtifactStore(str(tmp_path), resume=True)
    calls = []
    fn = lambda: (calls.append(1), {"x": torch.arange(3)})[1]
    a = store.run("stage1", "h1", fn, log)
    b = store.run("stage1", "h1", fn, log)
    assert len(calls) == 1 and torch.equal(a["x"], b["x"])
    store.run("stage1", "h2", fn, log)                 # configuration changed -> recomputed
    assert len(calls) == 2
    with log.stage("demo", bytes_moved=1e6):
        pass
    recs = [json.loads(l) for l in open(tmp_path / "run.jsonl")]
    assert any(r["event"] == "stage_skipped" for r in recs) and any(r.get("stage") == "demo" and "gb_per_s" in r for r in recs)
    with pytest.raises(NumericalFailure):
        check_finite("w", torch.tensor([1.0, float("nan")]))
    assert "lambda index 7" in describe_jerr(-7) and "capacity" in describe_jerr(-10012)


@pytest.mark.slow
def test_pipeline_small_cpu(tmp_path):
    cfg = C.load(None, ["io.out_dir=" + str(tmp_path), "device=cpu", "lasso.exclusion_reps=1", "rf.n_trees=[60]", "rf.proximity=false",
                        'gbm.reps=[{"max_depth": 2, "nrounds": 2}]', 'svm.kernels=["linear"]', "svm.weightfilt=[null]"])
    s = pipeline.run(cfg)
    assert s["lasso"]["confusion"] == [[21, 2], [1, 20]] and abs(s["lasso"]["auc"] - 0.9835) < 1e-3
    assert s["xgb1"]["n_genes"] == 6 and s["rf_60"]["oob_error"] < 0.2 and s["svm1"]["auc"] > 0.9
    for f in ("standouttable.csv", "consensus_rank.csv", "membership_table.csv", "Logistic_Lasso_RiskGroup_Associated_Genes.csv",
              "run.jsonl", "report/volcano.csv", "report/lasso_rep_curves.csv"):
        assert os.path.exists(tmp_path / f), f
    # second run resumes the lasso stage from its checkpoint
    pipeline.run(cfg)
    recs = [json.loads(l) for l in open(tmp_path / "run.jsonl")]
    assert any(r["event"] == "stage_skipped" and r["stage"] == "tasks/lasso" for r in recs)
    # ... but not for a different cohort of the same shape (the key used to hold only the shape)
    from consensusml_b200 import datasets
    co = datasets.load_target_aml()
    co.assay = co.assay.clone()
    co.assay[0, 0] += 1.0
    n_before = sum(r["event"] == "stage_skipped" for r in recs)
    pipeline.run(cfg, co)
    recs = [json.loads(l) for l in open(tmp_path / "run.jsonl")]
    assert sum(r["event"] == "stage_skipped" for r in recs) == n_before


def test_task_assignment_is_deterministic_and_balanced():
    from consensusml_b200.parallel.tasks import Task, assign
    tasks = [Task("lasso", 15.0, lambda: 0)] + [Task(f"rf{i}", 1.2, lambda: 0) for i in range(17)] + [Task(f"svm{i}", 1.0, lambda: 0) for i in range(4)]
    for world in (1, 2, 4, 8):
        a = assign(tasks, world)
        assert a == assign(tasks, world) and set(a.values()) <= set(range(world))
        load = [sum(t.cost for t in tasks if a[t.name] == r) for r in range(world)]
        assert max(load) <= max(15.0, sum(t.cost for t in tasks) / world + 1.2)       # LPT bound: optimum + one small task


def test_pipeline_two_ranks_gloo_matches_single_process(tmp_path):
    """The ensemble spread over two processes (gloo) gives the same standouttable as one process."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = os.path.join(root, "tests", "dist_pipeline_check.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29547", script, str(tmp_path)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env={**os.environ, "CML_DIST_TIMEOUT_S": "120"})
    assert r.returncode == 0 and "PIPELINE DIST OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_counts_to_cohort_cpu():
    from consensusml_b200.io.synthetic import nb_counts
    c, g = nb_counts(4000, 48, seed=2, n_de=250)
    train = np.arange(48) % 3 != 0
    co = pipeline.cohort_from_counts(torch.from_numpy(c), [f"ENSG{i:011d}.1" for i in range(4000)], [f"s{i}" for i in range(48)], g, train,
                                     C.Config())
    assert 20 < co.n_genes < 600 and co.n_samples == 48
    assert (co.row_data["p.adj.bh"] < 0.05).all() and (co.row_data["logFC"].abs() > 1).all()
    tr, te = co.split()
    assert tr.n_samples == 32 and te.n_samples == 16


def test_cli_config_and_lasso(tmp_path, capsys):
    from consensusml_b200.cli.main import main
    assert main(["config", "lasso.alpha=0.5"]) == 0
    assert json.loads(capsys.readouterr().out)["lasso"]["alpha"] == 0.5
    assert main(["lasso", "device=cpu", "io.out_dir=" + str(tmp_path)]) == 0
    out = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert out["genes"] == 13 and out["confusion"] == [[21, 2], [1, 20]]


def test_multiclass_one_vs_rest_cpu(tmp_path):
    """Pan-cancer style run (BASELINE.json config 3) on a small synthetic 3-class count matrix."""
    from consensusml_b200.io.synthetic import nb_counts_multiclass
    c, lab = nb_counts_multiclass(3000, (30, 28, 26), n_de_per_class=150, seed=4)
    train = np.arange(lab.size) % 3 != 0
    cfg = C.load(None, ["io.out_dir=" + str(tmp_path), "device=cpu", "lasso.exclusion_reps=1", "lasso.nfolds=4", "rf.n_trees=[40]",
                        "rf.proximity=false", 'gbm.reps=[{"max_depth": 2, "nrounds": 3}]', "svm.enabled=false", "mlp.enabled=false"])
    co = pipeline.multiclass_cohort_from_counts(torch.from_numpy(c), [f"ENSG{i:011d}.1" for i in range(3000)],
                                                [f"s{i}" for i in range(lab.size)], lab, train, cfg, max_genes_per_class=60)
    assert 30 <= co.n_genes <= 180
    s = pipeline.run_multiclass(cfg, co, class_names={0: "AML", 1: "NBL", 2: "WT"})
    assert s["classes"] == ["AML", "NBL", "WT"] and s["accuracy"] >= 0.85
    assert np.asarray(s["confusion"]).sum() == int((~train).sum())
    assert len(s["top"]) > 0 and os.path.exists(tmp_path / "consensus_rank_multiclass.csv")
    assert os.path.exists(tmp_path / "class_AML" / "standouttable.csv")


def test_cli_concat_then_run_on_counts(tmp_path, capsys):
    """Per-sample HTSeq files -> counts CSV -> normalise + DE prefilter + ensemble, all through the CLI."""
    import pandas as pd
    from consensusml_b200.cli.main import main
    from consensusml_b200.io.synthetic import nb_counts
    c, g = nb_counts(2500, 42, seed=5, n_de=200)
    files = []
    for j in range(42):
        f = tmp_path / f"S{j:03d}.htseq.counts"
        with open(f, "w") as fh:
            for i in range(2500):
                fh.write(f"ENSG{i:011d}.1\t{c[i, j]}\n")
            fh.write("__no_feature\t123\n")
        files.append(str(f))
    out = tmp_path / "out"
    assert main(["concat", "--files", str(tmp_path / "*.htseq.counts"), "io.out_dir=" + str(out)]) == 0
    info = json.loads(capsys.readouterr().out)
    assert info["genes"] == 2500 and info["samples"] == 42
    pd.DataFrame({"sample": [f"S{j:03d}" for j in range(42)], "group": np.where(g == 1, "High", "Low")}).to_csv(tmp_path / "groups.csv", index=False)
    rc = main(["run", "--counts", info["csv"], "--groups", str(tmp_path / "groups.csv"), "io.out_dir=" + str(out), "device=cpu",
               "lasso.exclusion_reps=1", "lasso.nfolds=4", "rf.n_trees=[30]", "rf.proximity=false", 'gbm.reps=[{"max_depth": 2, "nrounds": 2}]',
               "svm.enabled=false", "mlp.enabled=false"])
    assert rc == 0
    summary = json.loads(capsys.readouterr().out)
    assert summary["lasso"]["auc"] > 0.8 and os.path.exists(out / "standouttable.csv") and os.path.exists(out / "consensus_rank.csv")


def test_cli_normalize_and_de_on_counts_csv(tmp_path, capsys):
    """`normalize` writes the TMM log2-CPM matrix, `de` the limma-voom table in the reference's column layout."""
    import pandas as pd
    from consensusml_b200.cli.main import main
    from consensusml_b200.io.counts import write_matrix_csv
    from consensusml_b200.io.synthetic import nb_counts
    c, g = nb_counts(1500, 24, seed=9, n_de=120)
    genes, samples = [f"ENSG{i:011d}.2" for i in range(1500)], [f"S{j:02d}" for j in range(24)]
    write_matrix_csv(str(tmp_path / "counts.csv"), torch.from_numpy(c), genes, samples)
    pd.DataFrame({"sample": samples, "group": g}).to_csv(tmp_path / "groups.csv", index=False)
    assert main(["normalize", "--counts", str(tmp_path / "counts.csv"), "io.out_dir=" + str(tmp_path), "device=cpu"]) == 0
    info = json.loads(capsys.readouterr().out)
    norm = pd.read_csv(tmp_path / "TMMCPM_log2_Norm_Counts.csv", index_col=0)
    assert norm.shape == (info["genes_kept"], 24) and 0.5 < info["norm_factor_range"][0] <= 1 <= info["norm_factor_range"][1] < 2
    assert main(["de", "--counts", str(tmp_path / "counts.csv"), "--groups", str(tmp_path / "groups.csv"), "io.out_dir=" + str(tmp_path),
                 "device=cpu"]) == 0
    info = json.loads(capsys.readouterr().out)
    deg = pd.read_csv(info["csv"], index_col=0)
    assert list(deg.columns) == ["logFC", "AveExpr", "t", "P.Value", "adj.P.Val", "B"] and len(deg) == info["degs"] > 10
    assert (deg["adj.P.Val"] < 0.05).all() and (deg["logFC"].abs() > 1).all() and deg["P.Value"].is_monotonic_increasing


def test_failed_stage_is_reported_not_fatal_when_fail_fast_is_off(tmp_path, monkeypatch):
    from consensusml_b200.models import svm as S
    def boom(*a, **k):
        raise RuntimeError("injected fault")
    monkeypatch.setattr(S, "run_svm", boom)
    ov = ["io.out_dir=" + str(tmp_path), "device=cpu", "lasso.exclusion_reps=1", "rf.n_trees=[30]", "rf.proximity=false",
          'gbm.reps=[{"max_depth": 2, "nrounds": 2}]', 'svm.kernels=["linear"]', "svm.weightfilt=[null]", "checkpoint=false"]
    s = pipeline.run(C.load(None, ov + ["fail_fast=false"]))
    assert "injected fault" in s["failed_tasks"]["svm1"] and "svm1" not in s and s["lasso"]["confusion"] == [[21, 2], [1, 20]]
    assert len(s["consensus"]["top"]) == 50
    with pytest.raises(RuntimeError, match="injected fault"):
        pipeline.run(C.load(None, ov))


@pytest.mark.slow
def test_pipeline_on_the_kernel_path_under_emulation(tmp_path):
    """pipeline.run with every stage on the kernel code paths (lasso path kernel in latency mode, SMO, tree level kernels,
    metrics, MLP tail / Adam, replicate weights, stability selection), executed by the host execution model."""
    from consensusml_b200 import _native
    cfg = C.load(None, ["io.out_dir=" + str(tmp_path), "device=cpu", "lasso.exclusion_reps=1", "lasso.nfolds=4", "lasso.stability_boot=4",
                        "lasso.stability_chunk=2", "rf.n_trees=[60]", "rf.chunk=30", 'gbm.reps=[{"max_depth": 2, "nrounds": 2}]',
                        'svm.kernels=["linear"]', "svm.weightfilt=[null, 0.5]", "mlp.enabled=true", "mlp.epochs=15", "mlp.hidden=[8]",
                        "mlp.lr=[0.003]", "mlp.weight_decay=[0.0]", "checkpoint=false"])
    with _native.emulation(seed=1, sms=8):
        s = pipeline.run(cfg)
    assert "failed_tasks" not in s
    assert len(s["lasso"]["genes_rep1"]) == 13 and s["lasso"]["auc"] > 0.95           # 4-fold CV picks a neighbouring lambda
    assert s["xgb1"]["n_genes"] == 6 and abs(s["xgb1"]["train_error"] - 0.010753) < 1e-5 + 0.02
    assert s["rf_60"]["oob_error"] < 0.2 and s["svm1"]["auc"] > 0.9 and s["svm2"]["auc"] > 0.9 and s["mlp"]["auc"] > 0.85
    assert s["stability"]["stable_genes"] > 5 and len(s["consensus"]["top"]) == 50
    assert os.path.exists(tmp_path / "standouttable.csv")


def test_overlapped_stages_give_the_same_tables(tmp_path):
    """dist.task_streams > 1: the rank's tasks run on a pool of host threads (one CUDA stream each on a GPU); every task draws its
    randomness from its own identity, so the result files are those of the sequential run."""
    import pandas as pd
    ov = ["device=cpu", "lasso.exclusion_reps=2", "rf.n_trees=[60, 90]", "rf.chunk=30", "rf.proximity=false", "checkpoint=false",
          'gbm.reps=[{"max_depth": 2, "nrounds": 2}, {"max_depth": 3, "nrounds": 2}]', 'svm.kernels=["linear", "radial"]', "svm.weightfilt=[null]"]
    a = pipeline.run(C.load(None, ov + ["io.out_dir=" + str(tmp_path / "seq")]))
    b = pipeline.run(C.load(None, ov + ["io.out_dir=" + str(tmp_path / "par"), "dist.task_streams=3"]))
    assert a["lasso"] == b["lasso"] and a["consensus"]["top"] == b["consensus"]["top"]
    ta, tb = pd.read_csv(tmp_path / "seq" / "standouttable.csv"), pd.read_csv(tmp_path / "par" / "standouttable.csv")
    pd.testing.assert_frame_equal(ta, tb)
    assert b["schedule"]["rank_seconds"][0] > 0 and set(b["schedule"]["group_seconds"]) == set(a["schedule"]["group_seconds"])


def test_overlapped_task_runner_propagates_failures():
    import time as _t
    from consensusml_b200.parallel.tasks import Task, TaskError, run_tasks
    seen = []

    def ok(i):
        def f():
            _t.sleep(0.02)
            seen.append(i)
            return i
        return f

    def bad():
        raise ValueError("boom")
    tasks = [Task(f"t{i}", float(i), ok(i)) for i in range(6)] + [Task("bad", 10.0, bad)]
    res, rep = run_tasks(tasks, fail_fast=False, streams=3)
    assert [res[f"t{i}"] for i in range(6)] == list(range(6)) and isinstance(res["bad"], TaskError) and sorted(seen) == list(range(6))
    assert rep.rank_seconds[0] < sum(rep.seconds.values())            # the intervals overlapped
    with pytest.raises(ValueError, match="boom"):
        run_tasks(tasks, fail_fast=True, streams=3)
