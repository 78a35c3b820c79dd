"""Component-wise L2 gradient boosting of linear base learners - mboost ``glmboost`` as driven through mlr's
``regr.glmboost`` in Ryan/Feature_selection_GLMTrain_WT_TARGET.R:56-64 (Gaussian family, nu = 0.1, mstop = 100,
centred predictors), with that script's prefilter (between-cohort mean difference, :44-53) and its dichotomised
evaluation (:78-108).

Each boosting step needs the univariate least-squares fit of the residual on every column: ``X^T r`` (the same
product as the lasso gradient - the fused reduce-scatter path when sample-sharded) followed by an arg-max."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Sequence

import torch


@dataclass
class GLMBoostFit:
    intercept: float
    coef: torch.Tensor            # [p] on the original (uncentred) scale
    path: list                    # selected feature per step
    risk: list                    # mean squared residual per step

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return self.intercept + x.double() @ self.coef

    def selected(self) -> torch.Tensor:
        return torch.nonzero(self.coef != 0).flatten()


def glmboost(x: torch.Tensor, y: torch.Tensor, mstop: int = 100, nu: float = 0.1, center: bool = True) -> GLMBoostFit:
    xd, yd = x.double(), y.double()
    mean = xd.mean(0) if center else torch.zeros(xd.shape[1], dtype=torch.float64, device=x.device)
    xc = xd - mean
    ss = (xc * xc).sum(0)
    usable = ss > 1e-12
    offset = yd.mean()
    r = yd - offset
    coef = torch.zeros(xd.shape[1], dtype=torch.float64, device=x.device)
    path, risk = [], []
    for _ in range(mstop):
        g = xc.t() @ r                                         # X^T r
        score = torch.where(usable, g * g / ss.clamp_min(1e-12), torch.zeros_like(g))
        j = int(score.argmax())
        bj = g[j] / ss[j]
        coef[j] += nu * bj
        r = r - nu * bj * xc[:, j]
        path.append(j)
        risk.append(float((r * r).mean()))
    intercept = float(offset - (coef * mean).sum())
    return GLMBoostFit(intercept, coef, path, risk)


def mean_difference_filter(x: torch.Tensor, group: torch.Tensor, threshold: float = 0.05) -> torch.Tensor:
    """Columns whose between-group mean difference exceeds ``threshold`` in absolute value (RYN:44-53)."""
    g = group.bool()
    return torch.nonzero((x[g].double().mean(0) - x[~g].double().mean(0)).abs() > threshold).flatten()


def dichotomy_metrics(pred: torch.Tensor, truth: torch.Tensor, cutoff: float) -> Dict[str, float]:
    """Binary metrics after thresholding predicted and true survival times at ``cutoff`` (RYN:78-108: 331 days).
    'Positive' = below the cutoff (an early event)."""
    p, t = pred <= cutoff, truth <= cutoff
    tp, fp = float((p & t).sum()), float((p & ~t).sum())
    fn, tn = float((~p & t).sum()), float((~p & ~t).sum())
    d = lambda a, b: a / b if b else float("nan")
    return {"tpr": d(tp, tp + fn), "precision": d(tp, tp + fp), "specificity": d(tn, tn + fp), "fdr": d(fp, tp + fp),
            "mae": float((pred.double() - truth.double()).abs().mean())}
