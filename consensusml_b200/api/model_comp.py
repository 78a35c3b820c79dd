"""The reference's modelling toolkit ``scripts/model_comp.py`` (lines 6-170) on the B200 tree engine.

Same entry points and hyper-parameters; the three CPU libraries are replaced by
:class:`consensusml_b200.models.trees.GradientBoosting` (xgboost and sklearn modes) and
:class:`~consensusml_b200.models.trees.RandomForest`.  Data preparation keeps the reference's magic offsets
(gene columns are ``columns[84:-4]``) as overridable defaults."""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import pandas as pd
import torch

from ..models.trees import GradientBoosting, RandomForest
from ..ops.metrics import log_loss


def _dev(device):
    if device is not None:
        return torch.device(device)
    return torch.device("cuda" if torch.cuda.is_available() else "cpu")


def _t(a, device):
    return torch.as_tensor(np.asarray(a, dtype=np.float32), device=device)


def model_comp(X_train, X_test, y_train, y_test, device=None, verbose: bool = True):
    """Fit XGB(lr .01, depth 3, 700 trees), GB(lr .01, depth 4, log2 features, min leaf 4, 280 trees, subsample
    .25) and RF(300 trees, depth 3), average their class-1 probabilities, print the four log-losses and
    return ``(xgboost_model, random_forest_model)`` like the reference (model_comp.py:6-34).  The fitted
    gradient-boosting model and the losses are attached to the returned xgboost model as ``.extras``."""
    import time
    dev = _dev(device)
    xtr, xte, ytr, yte = _t(X_train, dev), _t(X_test, dev), _t(y_train, dev), _t(y_test, dev)

    def timed(fn):
        if dev.type == "cuda":
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = fn()
        if dev.type == "cuda":
            torch.cuda.synchronize()
        return out, time.perf_counter() - t0
    seconds = {}
    (xgb, p_xgb), seconds["xgboost"] = timed(lambda: (lambda m: (m, m.predict_proba(xte)))(
        GradientBoosting(n_estimators=700, learning_rate=0.01, max_depth=3, mode="xgboost", seed=8).fit(xtr, ytr)))
    (gb, p_gb), seconds["gradient_boost"] = timed(lambda: (lambda m: (m, m.predict_proba(xte)))(
        GradientBoosting(n_estimators=280, learning_rate=0.01, max_depth=4, max_features="log2", min_samples_leaf=4,
                         subsample=0.25, mode="sklearn", seed=8).fit(xtr, ytr)))
    (rf, p_rf), seconds["random_forest"] = timed(lambda: (lambda m: (m, m.predict_proba_mean(xte)))(
        RandomForest(n_trees=300, max_depth=3, seed=8).fit(xtr, ytr)))
    ens = (p_rf + p_gb + p_xgb) / 3
    losses = {"ensemble": log_loss(yte, ens), "gradient_boost": log_loss(yte, p_gb), "random_forest": log_loss(yte, p_rf),
              "xgboost": log_loss(yte, p_xgb)}
    if verbose:
        print("Ensemble Log Loss " + str(losses["ensemble"]))
        print("Gradient Boost Log Loss " + str(losses["gradient_boost"]))
        print("Random Forest Log Loss " + str(losses["random_forest"]))
        print("XGBoost Log Loss " + str(losses["xgboost"]))
    xgb.extras = {"gradient_boost_model": gb, "log_loss": losses, "ensemble_proba": ens, "seconds": seconds}
    return xgb, rf


def data_prep_columns(df: pd.DataFrame, var: str, gene_slice: Tuple[int, Optional[int]] = (84, -4), k: int = 1000):
    """Keep diagnostic ids 09A/03A/01A, drop ``Risk group == 'Unknown'``, add ``Low Risk``; per class take the
    ``k`` genes of highest (``var='Max'``) or lowest variance and return ``(frame, column positions)``
    (model_comp.py:37-80).  The per-class variance + top-k is the K12 reduction (ops/colvar.py)."""
    from ..ops.colvar import variance_topk
    reverse = var == "Max"
    df1 = df[df["Diagnostic ID"].isin(("09A", "03A", "01A"))]
    df2 = df1[df1["Risk group"] != "Unknown"].copy()
    df2["Low Risk"] = (df2["Risk group"] == "Low").astype(int)
    lo, hi = gene_slice
    gene_cols = df2.columns[lo:hi] if hi is not None else df2.columns[lo:]
    take: List[int] = []
    for cls in (1, 0):
        sub = df2.loc[df2["Low Risk"] == cls, gene_cols].to_numpy(dtype=np.float32)
        idx = variance_topk(torch.from_numpy(np.array(sub)), k, largest=reverse).tolist()
        for i in idx:
            pos = i + lo
            if pos not in take:
                take.append(pos)
    return df2, take


def _prep(data: pd.DataFrame, y: pd.Series):
    data = data.copy()
    data["label"] = y
    holdout = data.sample(frac=0.2, random_state=8)
    train = data.drop(holdout.index)
    X, yv = train.iloc[:, :-1], train["label"]
    from sklearn.model_selection import train_test_split
    X_train, X_test, y_train, y_test = train_test_split(X.values, yv, test_size=0.33, random_state=8)
    return X_train, X_test, y_train, y_test, holdout, X


def model_prep(df: pd.DataFrame, columns_to_take: Sequence[int]):
    """20 % holdout (``sample(frac=.2, random_state=8)``) then ``train_test_split(test_size=.33, random_state=8)``
    (model_comp.py:82-108)."""
    return _prep(df.iloc[:, list(columns_to_take)], df["Low Risk"])


def model_prep_loc(df: pd.DataFrame, columns_to_take: Sequence[str]):
    """Label-based variant (model_comp.py:110-121)."""
    return _prep(df.loc[:, list(columns_to_take)], df["Low Risk"])


def xgboost_tuner(X_train, X_test, y_train, y_test, n_estimators_num: Sequence[int], device=None, verbose: bool = True):
    """Log-loss at ``100 * n`` boosting rounds for each n (model_comp.py:123-170).  The reference fits five models
    from scratch; boosting is incremental, so one model with max(n)*100 rounds is evaluated at each prefix."""
    dev = _dev(device)
    xtr, xte, ytr, yte = _t(X_train, dev), _t(X_test, dev), _t(y_train, dev), _t(y_test, dev)
    trees = [100 * int(x) for x in n_estimators_num]
    model = GradientBoosting(n_estimators=max(trees), learning_rate=0.01, max_depth=6, mode="xgboost", seed=8).fit(xtr, ytr)
    out = {}
    for t in trees:
        out[t] = log_loss(yte, model.predict_proba(xte, n_trees=t))
        if verbose:
            print(f"XGB n_estimators: {t} log loss " + str(out[t]))
    return out
