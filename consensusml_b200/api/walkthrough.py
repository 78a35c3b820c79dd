"""The reference's Python end-to-end run, ``scripts/model_walkthrough.ipynb``, as one function.

The notebook builds the AML frame with ``clean.py``, then runs ``model_comp`` on three kinds of gene
subsets - per-class variance-selected (raw lines 1169-1201), three random draws of ~2000 genes (1244-1424) and all genes
(1569-1645) - fits an L1 logistic regression on all genes (1471-1476, 3886-3891), turns every model into a gene set with
"importance > 0" (1063-1064, 1488-1489, 1665-1666, 3914), intersects the sets (1796-1943, 4878-4991), writes the lasso gene
list (4830) and scores the 20 % holdout (1727-1739, 4805).  ``run_walkthrough`` does the same on the B200 tree engine and the
path solver; tree counts can be scaled down for quick runs."""
from __future__ import annotations

from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import pandas as pd
import torch

from .. import consensus as CS
from ..models.lasso import l1_logistic
from ..models.trees import GradientBoosting, RandomForest
from ..ops.metrics import log_loss
from . import model_comp as MC


def random_gene_columns(n_genes: int, k: int = 2000, offset: int = 84, rng: Optional[np.random.Generator] = None) -> list:
    """``np.random.choice(n_genes, k)`` WITH replacement, de-duplicated through a set, shifted by the clinical-column
    offset - the notebook's draw (raw lines 1245-1248), which therefore yields slightly fewer than ``k`` columns."""
    rng = rng or np.random.default_rng()
    cols = sorted(set(rng.integers(0, n_genes, k).tolist()))
    return [c + offset for c in cols]


def _fit_trees(X_train, X_test, y_train, y_test, device, scale: float, verbose: bool):
    """``model_comp`` with the tree counts multiplied by ``scale`` (1.0 = the reference's 700 / 280 / 300)."""
    if scale == 1.0:
        return MC.model_comp(X_train, X_test, y_train, y_test, device=device, verbose=verbose)
    dev = MC._dev(device)
    xtr, xte, ytr, yte = MC._t(X_train, dev), MC._t(X_test, dev), MC._t(y_train, dev), MC._t(y_test, dev)
    k = lambda n: max(2, int(round(n * scale)))
    xgb = GradientBoosting(n_estimators=k(700), learning_rate=0.01 / scale, max_depth=3, mode="xgboost", seed=8).fit(xtr, ytr)
    gb = GradientBoosting(n_estimators=k(280), learning_rate=0.01 / scale, max_depth=4, max_features="log2", min_samples_leaf=4,
                          subsample=0.25, mode="sklearn", seed=8).fit(xtr, ytr)
    rf = RandomForest(n_trees=k(300), max_depth=3, seed=8).fit(xtr, ytr)
    p_rf, p_gb, p_xgb = rf.predict_proba_mean(xte), gb.predict_proba(xte), xgb.predict_proba(xte)
    ens = (p_rf + p_gb + p_xgb) / 3
    xgb.extras = {"gradient_boost_model": gb, "ensemble_proba": ens,
                  "log_loss": {"ensemble": log_loss(yte, ens), "gradient_boost": log_loss(yte, p_gb),
                               "random_forest": log_loss(yte, p_rf), "xgboost": log_loss(yte, p_xgb)}}
    return xgb, rf


def run_walkthrough(df: pd.DataFrame, *, gene_slice: Tuple[int, Optional[int]] = (84, -4), k_var: int = 1000, k_random: int = 2000,
                    n_random: int = 3, seed: int = 0, C: float = 1.0, device=None, tree_scale: float = 1.0,
                    out_csv: Optional[str] = None, verbose: bool = False) -> Dict:
    """``df``: the merged clinical + assay frame produced by ``api.clean`` (genes in ``columns[gene_slice]``).

    Returns a dict with the log-losses of every experiment, the per-model gene sets of the all-genes experiment, their
    intersections and membership table, the L1-logistic gene list and the holdout log-losses."""
    dev = MC._dev(device)
    out: Dict = {"experiments": {}}
    frame, var_cols = MC.data_prep_columns(df, "Max", gene_slice=gene_slice, k=k_var)
    lo, hi = gene_slice
    n_cols = frame.shape[1]
    gene_pos = list(range(lo, n_cols + hi if hi is not None and hi < 0 else (hi or n_cols)))
    gene_names = [frame.columns[i] for i in gene_pos]

    def experiment(name, cols):
        Xtr, Xte, ytr, yte, holdout, X = MC.model_prep(frame, cols)
        xgb, rf = _fit_trees(Xtr, Xte, ytr, yte, dev, tree_scale, verbose)
        out["experiments"][name] = {"n_genes": len(cols), **xgb.extras["log_loss"]}
        return xgb, rf, holdout, X

    experiment("variance_selected", var_cols)                                   # 1. per-class top-variance genes
    rng = np.random.default_rng(seed)
    for r in range(n_random):                                                   # 2. random gene subsets
        experiment(f"random_{r + 1}", random_gene_columns(len(gene_pos), k_random, lo, rng))
    xgb, rf, holdout, X = experiment("all_genes", gene_pos)                     # 3. every gene

    # 4. L1 logistic regression on all genes, sklearn parametrisation, selection rule coef_ > 0
    Xtr, Xte, ytr, yte, _, _ = MC.model_prep(frame, gene_pos)
    lr = l1_logistic(MC._t(Xtr, dev), MC._t(ytr, dev), C=C, positive_only=True)
    coef = lr["coef"].detach().cpu().numpy()
    eta = MC._t(Xte, dev) @ lr["coef"].to(dev).float() + float(lr["intercept"])
    out["experiments"]["l1_logistic"] = {"n_genes": int((coef > 0).sum()), "log_loss": log_loss(MC._t(yte, dev), torch.sigmoid(eta))}

    # 5. gene sets (importance > 0) and their intersections
    imp = {"xgboost": pd.Series(xgb.importance_["gain"].cpu().numpy(), index=gene_names),
           "random_forest": pd.Series(rf.importance_.cpu().numpy(), index=gene_names),
           "lasso": pd.Series(coef, index=gene_names)}
    sets = CS.selection_sets(imp, positive_only=True)
    out["set_sizes"] = {k: len(v) for k, v in sets.items()}
    out["intersections"] = {k: sorted(v) for k, v in CS.intersections(sets).items()}
    out["membership"] = CS.membership_table(sets)
    out["lasso_genes"] = sorted(sets["lasso"])
    if out_csv:
        pd.DataFrame({"gene": out["lasso_genes"]}).to_csv(out_csv)                # lasso_genes_178.csv in the reference

    # 6. the 20 % holdout, scored with the all-genes models
    hx = MC._t(holdout.iloc[:, :-1].to_numpy(), dev)
    hy = MC._t(holdout["label"].to_numpy(), dev)
    out["holdout"] = {"xgboost": log_loss(hy, xgb.predict_proba(hx)),
                      "l1_logistic": log_loss(hy, torch.sigmoid(hx @ lr["coef"].to(dev).float() + float(lr["intercept"])))}
    return out
