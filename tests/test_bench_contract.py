"""bench.py's own logic (argument handling, the two timing arms, accuracy gates, JSON contract) exercised on the CPU oracles
through its ``--device cpu`` dry run, with one process and under torchrun with two gloo ranks; and the reference arm."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL = ["lasso.exclusion_reps=1", "rf.n_trees=[40]", "rf.chunk=20", "rf.proximity=false", 'gbm.reps=[{"max_depth": 2, "nrounds": 2}]',
         'svm.kernels=["linear"]', "svm.weightfilt=[null]"]
KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
        "config", "clocks", "e2e", "gpu_launches"}


def _line(out):
    return json.loads([l for l in out.splitlines() if l.startswith("{")][-1])


@pytest.mark.parametrize("nproc", [1, 2])
def test_bench_dry_run_prints_the_contract_line(nproc):
    args = [os.path.join(ROOT, "bench.py"), "--device", "cpu", "--gpus", str(nproc), "--steps", "1", "--warmup", "0", "--fits", "3", "--no-fused"] + SMALL
    if nproc == 1:
        cmd = [sys.executable] + args
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
               "--master-port", "29571"] + args
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env={**os.environ, "CML_DIST_TIMEOUT_S": "120"})
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    d = _line(r.stdout)
    assert KEYS <= set(d) and d["metric"] == "consensus_pipeline_wall_s" and d["higher_is_better"] is False and d["n_gpus"] == nproc
    assert "invalid" in d                                              # a dry run can never be mistaken for a measurement
    acc = d["accuracy"]
    assert acc["lasso_confusion_matches_reference"] and acc["lasso_genes_matching_reference_13"] == 13 and acc["lasso_auc_abs_err"] < 0.01
    assert d["e2e"]["h2d_bytes_per_step"] == 1984 * 137 * 4 and d["e2e"]["d2h_bytes_per_step"] > 0 and d["e2e"]["value"] > 0
    assert len(d["schedule"]["rank_seconds"]) == nproc and d["secondary"]["fits"]["lasso_path_fits_per_s"] > 0


def test_reference_arm_runs_the_unmodified_reference_or_says_why():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    d = _line(r.stdout)
    assert d["impl"] == "reference"
    assert isinstance(d["unavailable"], str) and d["unavailable"] and "value" not in d      # no number under the pipeline's metric name
    if "partial" not in d:
        return
    pr = d["partial"]
    assert pr["metric"] == "model_comp_wall_s" and pr["value"] > 0 and "substitutions" in pr["config"] and "value_scope" in pr
    ref_src = os.path.join(os.environ.get("CML_REFERENCE_DIR", "/root/reference"), "scripts", "model_comp.py")
    if os.path.exists(ref_src):                                        # the installed copy is byte-identical to the reference's file
        assert open(ref_src, "rb").read() == open(os.path.join(ROOT, "baseline", "_ref", "scripts", "model_comp.py"), "rb").read()
