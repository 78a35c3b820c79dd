"""The three side analyses as workflows (api/side_analyses.py) on synthetic stand-ins of their inputs: every step of the
scripts is exercised and the files they write come out in the reference's formats."""
import os

import numpy as np
import pandas as pd
import pytest

from consensusml_b200.api import side_analyses as SA


def _aml_counts(n_pat=60, n_deg=140, seed=0):
    rng = np.random.default_rng(seed)
    codes = [f"PA{chr(65 + i // 26)}{chr(65 + i % 26)}XX" for i in range(n_pat)]
    risk = rng.choice(["Low", "Standard", "High", "Unknown"], n_pat, p=[0.4, 0.3, 0.25, 0.05])
    clinical = pd.DataFrame({"TARGET.USI": [f"TARGET-20-{c}" for c in codes], "Risk.group": risk})
    clinical = pd.concat([clinical, clinical.iloc[:3]], ignore_index=True)                 # duplicated clinical rows
    cols = [f"TARGET.20.{c}.09A.01R" for c in codes] + [f"TARGET.20.{codes[0]}.03A.01R", "TARGET.30.PAZZZZ.01A.01R"]   # a second aliquot, a non-AML sample
    genes = [f"ENSG{i:011d}.{i % 5 + 1}" for i in range(n_deg + 40)] + ["__no_feature", "__ambiguous"]
    mu = np.exp(rng.normal(5, 1.2, len(genes)))[:, None] * np.ones((1, len(cols)))
    low = np.array([risk[codes.index(c.split(".")[2])] == "Low" if c.split(".")[2] in codes else False for c in cols])
    mu[:12, low] *= 3.0
    counts = pd.DataFrame(rng.negative_binomial(4, 4 / (4 + mu)), index=genes, columns=cols)
    deg_ids = genes[:n_deg] + ["__no_feature"]
    return counts, deg_ids, clinical, risk, codes


def test_mlseq_prepare_follows_the_script():
    counts, deg_ids, clinical, risk, codes = _aml_counts()
    prep = SA.mlseq_prepare(counts, deg_ids, clinical)
    known = int((risk != "Unknown").sum())
    assert prep["n_samples"] == known                                  # one column per patient with a known risk group
    assert prep["test"].shape[1] == int(np.ceil(0.3 * known)) and prep["train"].shape[1] == known - prep["test"].shape[1]
    assert len(prep["genes"]) == 100 and all(g in deg_ids and "__" not in g for g in prep["genes"])
    assert int(prep["train"].min().min()) >= 1                          # the +1 pseudo-count
    assert set(prep["y_train"]) <= {"H", "L"}
    var = counts.loc[[g for g in deg_ids if "__" not in g]].iloc[:, : len(codes)].var(axis=1)
    assert prep["genes"][0] in var.sort_values(ascending=False).index[:5]
    again = SA.mlseq_prepare(counts, deg_ids, clinical)
    assert np.array_equal(again["test_index"], prep["test_index"])      # seeded split


def test_mlseq_zoo_writes_the_gene_lists(tmp_path):
    counts, deg_ids, clinical, _, _ = _aml_counts(seed=1)
    prep = SA.mlseq_prepare(counts, deg_ids, clinical, n_genes=40)
    out = SA.mlseq_zoo(prep, str(tmp_path), methods={"PLDA": None, "PLDA2": None, "voomNSC": None, "pam": "deseq-vst", "svmRadial": "deseq-rlog"},
                       folds=3, repeats=1, tune_length=3)
    for m, rec in out["methods"].items():
        assert rec["test"]["accuracy"] >= 0.7, (m, rec["test"])
        assert 0 <= rec["test"]["sensitivity"] <= 1 and len(rec["selected_genes"]) >= 1
    for f in ("DEG_voomnsc.csv", "DEG_plda.csv", "DEG_plda2.csv"):
        lines = open(tmp_path / f).read().splitlines()
        assert lines[0] == '"","x"' and lines[1].startswith('"1","ENSG') and "." not in lines[1].split(",")[1]   # version-stripped ids


def test_wilms_glmboost_workflow(tmp_path):
    rng = np.random.default_rng(2)
    n, p = 120, 300
    x = rng.normal(size=(n, p))
    efs = np.clip(900 + 400 * x[:, 0] - 300 * x[:, 1] + 200 * x[:, 2] + rng.normal(scale=100, size=n), 5, None)
    x[:, 50:] *= 0.01                                                   # genes without a between-cohort difference
    df = pd.DataFrame({f"clin{i}": rng.integers(0, 3, n) for i in range(40)})
    df["Event.Free.Survival.Time.in.Days"] = efs
    df = pd.concat([df, pd.DataFrame(x, columns=[f"ENSG{i:011d}.1" for i in range(p)])], axis=1)
    out = SA.wilms_glmboost(df, split_days=float(np.median(efs)), cutoff=float(np.median(efs)), out_dir=str(tmp_path))
    assert out["train_equals_test"] and out["n_train"] == out["n_test"] == 60
    assert 3 <= out["n_genes_after_filter"] < p
    c = out["counts"]
    assert c["tp"] + c["fn"] == c["cp"] and c["fp"] + c["tn"] == c["cn"] and c["cp"] + c["cn"] == 60
    feats = pd.read_csv(tmp_path / "feature_list.csv")
    assert list(feats.columns) == ["gene", "coeff"] and {"ENSG00000000000.1", "ENSG00000000001.1"} <= set(feats["gene"])
    assert feats.set_index("gene").loc["ENSG00000000000.1", "coeff"] > 0 > feats.set_index("gene").loc["ENSG00000000001.1", "coeff"]
    honest = SA.wilms_glmboost(df, split_days=float(np.median(efs)), cutoff=float(np.median(efs)), honest_test=True)
    assert not honest["train_equals_test"] and honest["metrics"]["mae"] >= out["metrics"]["mae"]     # the script's number is optimistic


def test_km_whatever_three_aims(tmp_path):
    rng = np.random.default_rng(3)
    n = 200
    group = rng.integers(0, 3, n)
    proto = rng.random((3, 25)) < 0.3
    feats = pd.DataFrame((rng.random((n, 25)) < np.where(proto[group], 0.85, 0.05)).astype(int), columns=[f"mut{i}" for i in range(25)])
    risk = (rng.random(n) < np.array([0.15, 0.5, 0.85])[group]).astype(int)
    m, p = 130, 400
    y = rng.integers(0, 2, m)
    x = rng.normal(size=(m, p))
    x[:, :6] += 1.5 * y[:, None]
    x[:, 6:12] -= 1.5 * y[:, None]
    expr = pd.DataFrame(x, columns=[f"ENSG{i:011d}.2" for i in range(p)])
    flag = (np.arange(m) % 3 == 0).astype(int)                          # the notebook trains on the smaller, flagged set
    out = SA.km_whatever(feats, risk, expression=expr, expr_risk=y, train_flag=flag, ks=(6, 4, 3, 2), nstart=5, out_dir=str(tmp_path))
    km = out["kmeans"]
    assert list(km["k"]) == [6, 4, 3, 2] and km["tot_withinss"].is_monotonic_increasing
    assert km.set_index("k").loc[3, "deviance"] < km.set_index("k").loc[2, "deviance"] + 1e-9
    assert out["svm"]["auc"] > 0.9 and out["svm"]["n_train"] == int(flag.sum())
    hi4 = open(tmp_path / "output" / "High4.csv").read().splitlines()
    assert hi4[0] == '"","x"' and len(hi4) - 1 == out["svm"]["tail_sizes"][0.25][0] == 100
    assert {f"ENSG{i:011d}.2" for i in range(6)} <= {l.split(",")[1].strip('"') for l in hi4[1:]}
    assert len(open(tmp_path / "output" / "low1000.csv").read().splitlines()) - 1 == out["svm"]["tail_sizes"][0.001][1]
    assert sorted(os.listdir(tmp_path / "output")) == sorted([f"{a}{b}.csv" for a in ("High", "low") for b in ("4", "10", "20", "100", "1000")])
    assert out["qda"]["test_accuracy"] > 0.8


def test_cli_entry_points(tmp_path, capsys):
    import json
    from consensusml_b200.cli.main import main
    rng = np.random.default_rng(4)
    n, p = 80, 60
    x = rng.normal(size=(n, p))
    df = pd.DataFrame({f"clin{i}": rng.integers(0, 3, n) for i in range(40)})
    df["Event.Free.Survival.Time.in.Days"] = np.clip(500 + 300 * x[:, 0] + rng.normal(scale=50, size=n), 5, None)
    df = pd.concat([df, pd.DataFrame(x, columns=[f"g{i}" for i in range(p)])], axis=1)
    df.to_csv(tmp_path / "wt.csv", index=False)
    assert main(["glmboost", "--table", str(tmp_path / "wt.csv"), f"io.out_dir={tmp_path}", "device=cpu"]) == 0
    out = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert out["n_train"] == 40 and (tmp_path / "feature_list.csv").exists()
    y = rng.integers(0, 2, 90)
    e = pd.DataFrame(rng.normal(size=(90, 50)) + 1.2 * y[:, None] * (np.arange(50) < 5), columns=[f"g{i}" for i in range(50)])
    e["risk"], e["train"] = y, (np.arange(90) % 2)
    e.to_csv(tmp_path / "expr.csv", index=False)
    assert main(["kmsvm", "--expr", str(tmp_path / "expr.csv"), f"io.out_dir={tmp_path}", "device=cpu"]) == 0
    out = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert out["svm"]["auc"] > 0.8 and (tmp_path / "output" / "High4.csv").exists()
    counts, deg_ids, clinical, _, _ = _aml_counts(n_pat=40, n_deg=50, seed=6)
    counts.to_csv(tmp_path / "counts.csv")
    pd.DataFrame({"X": deg_ids}).to_csv(tmp_path / "degs.csv", index=False)
    clinical.to_csv(tmp_path / "clin.csv", index=False)
    SA.MLSEQ_METHODS_BACKUP = dict(SA.MLSEQ_METHODS)
    try:
        SA.MLSEQ_METHODS.clear()
        SA.MLSEQ_METHODS.update({"PLDA": None, "voomNSC": None})       # two quick members of the zoo
        import consensusml_b200.models.discriminant as DC
        orig = DC.classify
        DC.classify = lambda *a, **k: orig(*a, **{**k, "repeats": 1, "folds": 3})
        assert main(["mlseq", "--counts", str(tmp_path / "counts.csv"), "--degs", str(tmp_path / "degs.csv"), "--clinical-csv", str(tmp_path / "clin.csv"),
                     f"io.out_dir={tmp_path}", "device=cpu"]) == 0
    finally:
        DC.classify = orig
        SA.MLSEQ_METHODS.clear()
        SA.MLSEQ_METHODS.update(SA.MLSEQ_METHODS_BACKUP)
    assert (tmp_path / "DEG_plda.csv").exists() and (tmp_path / "DEG_voomnsc.csv").exists()


def test_phenotype_matrix_on_the_reference_workbook():
    """The real TARGET-AML clinical sheet (993 patients) through io.xlsx and the notebook's recoding, then its k-means for k = 4, 2."""
    from conftest import REFERENCE_DIR
    path = os.path.join(REFERENCE_DIR, "BuieRProj", "data", "TARGET_AML_ClinicalData_20160714.xlsx")
    if not os.path.exists(path):
        pytest.skip("reference tree not mounted")
    from consensusml_b200.io import xlsx as X
    clin = X.read_xlsx(path, keep_default_na=False)
    m = SA.phenotype_matrix(clin)
    assert m.shape == (993, 25) and set(np.unique(m.to_numpy())) <= {0.0, 1.0}
    assert {"t(8;21)", "inv(16)", "MLL", "FLT3/ITD positive?", "NPM mutation", "CEBPA mutation", "c-Kit Mutation Exon 17"} <= set(m.columns)
    assert 100 < m["t(8;21)"].sum() < 200 and 80 < m["inv(16)"].sum() < 200         # core-binding-factor AML: roughly a quarter of paediatric cases
    risk = (clin["Risk group"].astype(str) != "Low").astype(int).to_numpy()
    out = SA.km_whatever(m, risk, ks=(4, 2), nstart=5)
    km = out["kmeans"].set_index("k")
    assert km.loc[4, "tot_withinss"] < km.loc[2, "tot_withinss"] and km.loc[4, "deviance"] < km.loc[2, "deviance"] + 1e-9
    # the low-risk definition of the protocol IS cytogenetic (t(8;21), inv(16), NPM / CEBPA without FLT3-ITD): clusters must carry signal
    null_dev = -2 * (risk.sum() * np.log(risk.mean()) + (len(risk) - risk.sum()) * np.log(1 - risk.mean()))
    assert km.loc[4, "deviance"] < 0.8 * null_dev
