"""The reference's four archived notebooks (``scripts/archives/*.ipynb``) as functions.

They are the drafts that led to ``model_comp.py`` / ``model_walkthrough.ipynb``; what they compute that the final
notebook does not:

  david_clean     per-project clinical frames written to ``Clinical_Data/{AML,NBL,WT}_dataframe.csv``, value-count /
                  crosstab summaries, and a first raw-count merge keyed on ``barcode[:16]``       -> clinical_frames,
                                                                                                    raw_count_merge
  david_Eda       the 2000 genes of highest variance over ALL AML samples (``columns[84:-3]``)    -> eda_top_variance
  analysis_model  the inline prototype of ``model_comp``: diagnostic ids 09A / 03A only, per-class top-1000 variance
                  genes, 300 xgboost trees, top-100 genes by xgboost importance                  -> analysis_model
  model_testing   ``model_prep_lasso`` (since removed from model_comp.py), L1 logistic on all genes, then the cascade
                  lasso genes -> trees -> genes with importance > 0 -> trees again -> tuner      -> model_prep_lasso,
                                                                                                    lasso_cascade
The variance ranking is the K12 reduction (``ops.colvar``), the trees are the B200 tree engine, the L1 logistic fit is the
path solver in sklearn parametrisation."""
from __future__ import annotations

import os
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import pandas as pd
import torch

from ..models.lasso import l1_logistic
from ..models.trees import GradientBoosting, RandomForest
from ..ops.metrics import log_loss
from . import clean as CL
from . import model_comp as MC


def clinical_frames(manifest_df: pd.DataFrame, clinical: Dict[str, pd.DataFrame], out_dir: Optional[str] = None) -> Dict[str, pd.DataFrame]:
    """``clinical``: ``{"AML": sheet, "NBL": sheet, "WT": sheet}``; each is merged with the manifest rows of ``TARGET-<key>``
    and, with ``out_dir``, written as ``<key>_dataframe.csv`` (david_clean: the ``manifest_clinical_merge`` x3 and the
    three ``to_csv`` cells)."""
    out = {k: CL.manifest_clinical_merge(manifest_df, v, f"TARGET-{k}") for k, v in clinical.items()}
    if out_dir:
        os.makedirs(out_dir, exist_ok=True)
        for k, v in out.items():
            v.to_csv(os.path.join(out_dir, f"{k}_dataframe.csv"))
    return out


def summaries(frames: Dict[str, pd.DataFrame], manifest_df: pd.DataFrame) -> Dict[str, pd.Series]:
    """The value-count cells of david_clean: samples per project, and for every categorical column the notebook looks at
    (NBL: Histology, MKI, Ethnicity, MYCN status, Grade; AML: Risk group, First Event; WT: Stage, histology class) the
    level counts, where the column exists."""
    look = {"NBL": ["Histology", "MKI", "Ethnicity", "MYCN status", "Grade"], "AML": ["Risk group", "First Event"],
            "WT": ["Stage", "Histology Classification of Primary Tumor"]}
    out = {"project": manifest_df["project.project_id"].value_counts()}
    for k, cols in look.items():
        for c in cols:
            if k in frames and c in frames[k].columns:
                out[f"{k}.{c}"] = frames[k][c].value_counts()
    return out


def raw_count_merge(assay_df: pd.DataFrame, clinical_df: pd.DataFrame) -> pd.DataFrame:
    """david_clean's first attempt at joining expression to clinical rows, on the raw HTSeq matrix: transpose, split the
    column name into ``TARGET USI`` = first 16 characters and ``2nd ID`` = the rest, rewrite the clinical USI with dots and
    left-join on it.  (The final ``assay_clinical_merge`` joins on the whole barcode instead, which keeps replicate
    aliquots apart; this one returns one row per (patient, aliquot) pair.)  ``assay_df``: genes x samples, first column = gene ids."""
    t = assay_df.set_index(assay_df.columns[0]).T
    t.insert(0, "2nd ID", [str(i)[16:] for i in t.index])
    t.insert(0, "TARGET USI", [str(i)[:16] for i in t.index])
    clin = clinical_df.copy()
    clin["TARGET USI"] = clin["TARGET USI"].astype(str).str.replace("-", ".", regex=False)
    return clin.merge(t, how="left", on="TARGET USI")


def eda_top_variance(df: pd.DataFrame, k: int = 2000, gene_slice: Tuple[int, Optional[int]] = (84, -3)) -> list:
    """Column positions (in ``df``) of the ``k`` genes of highest variance over all rows, highest first (david_Eda: ``var``
    over ``columns[84:-3]``, sorted, ``[:2000]``, ``+ 84``)."""
    from ..ops.colvar import variance_topk
    lo, hi = gene_slice
    x = df.iloc[:, lo:hi].to_numpy(dtype=np.float32)
    idx = variance_topk(torch.from_numpy(np.ascontiguousarray(x)), min(k, x.shape[1]), largest=True).tolist()
    return [i + lo for i in idx]


def model_prep_lasso(df: pd.DataFrame, columns_to_take: Sequence[str]):
    """The removed helper model_testing.ipynb still calls: named gene columns -> 20 % holdout -> 67 / 33 split, as
    ``model_prep_loc`` (its successor in model_comp.py:110-121)."""
    return MC.model_prep_loc(df, columns_to_take)


def _three_models(Xtr, Xte, ytr, yte, dev, n_xgb: int, scale: float):
    xtr, xte, ytr_t, yte_t = MC._t(Xtr, dev), MC._t(Xte, dev), MC._t(ytr, dev), MC._t(yte, dev)
    k = lambda n: max(2, int(round(n * scale)))
    lr = 0.01 / scale
    xgb = GradientBoosting(n_estimators=k(n_xgb), learning_rate=lr, max_depth=3, mode="xgboost", seed=8).fit(xtr, ytr_t)
    gb = GradientBoosting(n_estimators=k(280), learning_rate=lr, max_depth=4, max_features="log2", min_samples_leaf=4, subsample=0.25,
                          mode="sklearn", seed=8).fit(xtr, ytr_t)
    rf = RandomForest(n_trees=k(300), max_depth=3, seed=8).fit(xtr, ytr_t)
    p_rf, p_gb, p_xgb = rf.predict_proba_mean(xte), gb.predict_proba(xte), xgb.predict_proba(xte)
    ens = (p_rf + p_gb + p_xgb) / 3
    losses = {"ensemble": log_loss(yte_t, ens), "gradient_boost": log_loss(yte_t, p_gb), "random_forest": log_loss(yte_t, p_rf),
              "xgboost": log_loss(yte_t, p_xgb)}
    return xgb, gb, rf, losses


def analysis_model(df: pd.DataFrame, *, gene_slice: Tuple[int, Optional[int]] = (84, -4), k: int = 1000, top: int = 100, device=None,
                   tree_scale: float = 1.0) -> Dict:
    """analysis_model.ipynb: rows with diagnostic id 09A / 03A (the final code also keeps 01A), ``Risk group != 'Unknown'``,
    ``Low Risk`` label; the union of the per-class ``k`` highest-variance genes; 20 % holdout + 67 / 33 split; xgboost with
    300 trees (700 in the final ``model_comp``), the same GB and RF; mean-probability ensemble; the ``top`` genes by xgboost
    importance.  ``tree_scale`` < 1 shrinks the tree counts (learning rate scaled up) for quick runs."""
    from ..ops.colvar import variance_topk
    dev = MC._dev(device)
    d = df[df["Diagnostic ID"].isin(("09A", "03A"))]
    d = d[d["Risk group"] != "Unknown"].copy()
    d["Low Risk"] = (d["Risk group"] == "Low").astype(int)
    lo, hi = gene_slice
    genes = d.columns[lo:hi]
    take: list = []
    for cls in (1, 0):
        sub = np.ascontiguousarray(d.loc[d["Low Risk"] == cls, genes].to_numpy(dtype=np.float32))
        for i in variance_topk(torch.from_numpy(sub), min(k, len(genes)), largest=True).tolist():
            if i + lo not in take:
                take.append(i + lo)
    Xtr, Xte, ytr, yte, holdout, X = MC.model_prep(d, take)
    xgb, gb, rf, losses = _three_models(Xtr, Xte, ytr, yte, dev, 300, tree_scale)
    imp = xgb.importance_["gain"].detach().cpu().numpy()
    order = np.argsort(-imp, kind="stable")[:top]
    return {"n_rows": int(len(d)), "columns": take, "log_loss": losses, "top_genes": [X.columns[i] for i in order],
            "models": {"xgboost": xgb, "gradient_boost": gb, "random_forest": rf}, "holdout": holdout}


def lasso_cascade(df: pd.DataFrame, *, gene_slice: Tuple[int, Optional[int]] = (84, -4), C: float = 1.0, device=None, tree_scale: float = 1.0,
                  tuner_rounds: Sequence[int] = (6, 7, 8, 9, 10)) -> Dict:
    """model_testing.ipynb after its ``model_comp`` experiments: L1 logistic regression on every gene (sklearn ``penalty='l1'``,
    ``C=1``), ``important_genes`` = coefficients > 0; the three tree models on those genes; the genes either forest gives
    positive importance; the models again on each of those two subsets; finally the xgboost tree-count sweep on the lasso
    genes (the notebook's conclusion: 700 trees).  ``df`` is the frame ``data_prep_columns`` returns (has ``Low Risk``)."""
    dev = MC._dev(device)
    lo, hi = gene_slice
    lasso_columns = df.columns[lo:hi]
    Xtr, Xte, ytr, yte, _, _ = model_prep_lasso(df, lasso_columns)
    lr = l1_logistic(MC._t(Xtr, dev), MC._t(ytr, dev), C=C, positive_only=True)
    coef = lr["coef"].detach().cpu().numpy()
    eta = MC._t(Xte, dev) @ lr["coef"].to(dev).float() + float(lr["intercept"])
    out: Dict = {"l1_log_loss": log_loss(MC._t(yte, dev), torch.sigmoid(eta))}
    important = lasso_columns[coef > 0]
    out["important_genes"] = list(important)
    if len(important) == 0:
        return out

    def fit_on(cols):
        a, b, c, d_, _, _ = MC.model_prep_loc(df, cols)
        xgb, _, rf, losses = _three_models(a, b, c, d_, dev, 700, tree_scale)
        return xgb, rf, losses

    xg, rf, out["on_lasso_genes"] = fit_on(important)
    xg_imp = important[xg.importance_["gain"].detach().cpu().numpy() > 0]
    rf_imp = important[rf.importance_.detach().cpu().numpy() > 0]
    out["xg_important"], out["rf_important"] = list(xg_imp), list(rf_imp)
    if len(rf_imp):
        out["on_rf_important"] = fit_on(rf_imp)[2]
    if len(xg_imp):
        out["on_xg_important"] = fit_on(xg_imp)[2]
    if tuner_rounds:
        a, b, c, d_, _, _ = MC.model_prep_loc(df, important)
        if tree_scale == 1.0:
            out["tuner"] = MC.xgboost_tuner(a, b, c, d_, list(tuner_rounds), device=dev, verbose=False)
        else:                                   # shortened sweep: same prefixes of one boosted model, fewer rounds per unit
            unit = max(1, int(round(100 * tree_scale)))
            model = GradientBoosting(n_estimators=unit * max(tuner_rounds), learning_rate=0.01 / tree_scale, max_depth=6, mode="xgboost",
                                     seed=8).fit(MC._t(a, dev), MC._t(c, dev))
            out["tuner"] = {100 * int(n): log_loss(MC._t(d_, dev), model.predict_proba(MC._t(b, dev), n_trees=unit * int(n))) for n in tuner_rounds}
    return out
