"""The archived notebooks' flows (api/archives.py) on a synthetic clinical + assay frame."""
import numpy as np
import pandas as pd

from consensusml_b200.api import archives as A
from consensusml_b200.api import model_comp as MC


def _frame(n=150, genes=200, seed=3):
    rng = np.random.default_rng(seed)
    df = pd.DataFrame({f"c{i}": rng.integers(0, 4, n) for i in range(84)}).rename(columns={"c1": "Diagnostic ID", "c2": "Risk group"})
    df["Diagnostic ID"] = rng.choice(["09A", "03A", "01A"], n, p=[0.6, 0.3, 0.1])
    risk = rng.choice(["Low", "Standard", "High", "Unknown"], n, p=[0.42, 0.28, 0.25, 0.05])
    df["Risk group"] = risk
    x = rng.normal(size=(n, genes)).astype(np.float32)
    x[:, :10] = 1.5 * x[:, :10] + (risk == "Low")[:, None] * 2.7   # informative AND above-average within-class variance
    x[:, 10:20] *= 3.0                                                 # high variance, uninformative
    g = pd.DataFrame(x, columns=[f"ENSG{i:05d}" for i in range(genes)])
    tail = pd.DataFrame({f"t{i}": rng.integers(0, 3, n) for i in range(3)})
    return pd.concat([df, g, tail], axis=1), x


def test_eda_top_variance_matches_pandas():
    frame, x = _frame()
    cols = A.eda_top_variance(frame, k=25)
    want = np.argsort(-frame.iloc[:, 84:-3].var(axis=0).to_numpy(), kind="stable")[:25] + 84
    assert cols == want.tolist()
    assert set(range(84 + 10, 84 + 20)) <= set(cols)


def test_clinical_frames_and_raw_merge(tmp_path):
    man = pd.DataFrame({"project.project_id": ["TARGET-AML", "TARGET-AML", "TARGET-NBL", "TARGET-WT"],
                        "entity_submitter_id": ["TARGET-20-PAAAAA-09A-01R", "TARGET-20-PABBBB-03A-01R", "TARGET-30-PACCCC-01A-01R", "TARGET-50-PADDDD-01A-01R"]})
    clin = {"AML": pd.DataFrame({"TARGET USI": ["TARGET-20-PAAAAA", "TARGET-20-PABBBB"], "Risk group": ["Low", "High"], "First Event": ["None", "Relapse"]}),
            "NBL": pd.DataFrame({"TARGET USI": ["TARGET-30-PACCCC"], "Histology": ["Unfavorable"], "MKI": ["Low"]}),
            "WT": pd.DataFrame({"TARGET USI": ["TARGET-50-PADDDD"], "Stage": ["III"]})}
    frames = A.clinical_frames(man, clin, str(tmp_path))
    assert {f.name for f in tmp_path.iterdir()} == {"AML_dataframe.csv", "NBL_dataframe.csv", "WT_dataframe.csv"}
    assert len(frames["AML"]) == 2 and frames["NBL"]["entity_submitter_id"].iloc[0].endswith("01A-01R")
    s = A.summaries(frames, man)
    assert s["project"]["TARGET-AML"] == 2 and s["AML.Risk group"]["Low"] == 1 and s["WT.Stage"]["III"] == 1 and "NBL.Grade" not in s
    assay = pd.DataFrame({"Genes": ["g1", "g2"], "TARGET.20.PAAAAA.09A.01R": [5, 0], "TARGET.20.PAAAAA.09A.02R": [7, 1], "TARGET.20.PABBBB.03A.01R": [1, 2]})
    merged = A.raw_count_merge(assay, frames["AML"])
    assert len(merged) == 3                                            # PAAAAA has two aliquots
    row = merged[merged["2nd ID"] == ".09A.02R"].iloc[0]
    assert row["g1"] == 7 and row["Risk group"] == "Low"


def test_analysis_model_prototype():
    frame, _ = _frame()
    out = A.analysis_model(frame, k=30, top=15, device="cpu", tree_scale=0.04)
    keep = frame["Diagnostic ID"].isin(("09A", "03A")) & (frame["Risk group"] != "Unknown")
    assert out["n_rows"] == int(keep.sum())                            # 01A rows are NOT kept by the prototype
    assert 30 <= len(out["columns"]) <= 60 and len(set(out["columns"])) == len(out["columns"])
    assert out["log_loss"]["ensemble"] < 0.66 and len(out["top_genes"]) == 15
    informative = {f"ENSG{i:05d}" for i in range(10)}
    assert len(informative & set(out["top_genes"])) >= 4


def test_lasso_cascade():
    frame, _ = _frame()
    df, _ = MC.data_prep_columns(frame, "Max", k=20)
    out = A.lasso_cascade(df, device="cpu", tree_scale=0.03, tuner_rounds=(1, 2, 3))
    informative = {f"ENSG{i:05d}" for i in range(10)}
    assert len(informative & set(out["important_genes"])) >= 5
    assert set(out["xg_important"]) <= set(out["important_genes"]) and set(out["rf_important"]) <= set(out["important_genes"])
    assert out["on_lasso_genes"]["ensemble"] < 0.66 and out["l1_log_loss"] < 0.69
    assert sorted(out["tuner"]) == [100, 200, 300] and all(np.isfinite(v) for v in out["tuner"].values())
    a = A.model_prep_lasso(df, ["ENSG00000", "ENSG00001"])
    b = MC.model_prep_loc(df, ["ENSG00000", "ENSG00001"])
    assert np.array_equal(a[0], b[0]) and a[4].equals(b[4])
