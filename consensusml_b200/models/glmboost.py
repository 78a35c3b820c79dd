"""Component-wise L2 gradient boosting of linear base learners - mboost ``glmboost`` as driven through mlr's
``regr.glmboost`` in Ryan/Feature_selection_GLMTrain_WT_TARGET.R:56-64 (Gaussian family, nu = 0.1, mstop = 100,
centred predictors), with that script's prefilter (between-cohort mean difference, :44-53) and its dichotomised
evaluation (:78-108).

Each boosting step needs the univariate least-squares fit of the residual on every column (``X^T r``) followed by an
arg-max and a rank-one residual update.  On CUDA all ``mstop`` steps are enqueued by one call of ``cml_glmboost``
(csrc/boost.cu: a warp-per-column scoring kernel and a one-CTA update kernel per step, stream-ordered, no host round
trip); CPU tensors run the float64 torch restatement below, which is also the oracle."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Dict, Sequence

import torch

from .. import _native

_I, _D, _P = ctypes.c_int, ctypes.c_double, ctypes.c_void_p
_native.register_signature("cml_glmboost", [_P, _I, _I, _I, _P, _P, _P, _I, _D, _P, _P, _P, _P])


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
    if _native.on_device(x):
        n, p = xd.shape
        n_pad = (n + 31) // 32 * 32
        xct = torch.zeros(p, n_pad, dtype=torch.float32, device=x.device)
        xct[:, :n] = xc.t().float()
        ssd = (xct.double() ** 2).sum(1).contiguous()          # sums of squares of the columns the kernel actually reads
        rp = torch.zeros(n_pad, dtype=torch.float64, device=x.device)
        rp[:n] = r
        best = torch.zeros(mstop, dtype=torch.int64, device=x.device)
        path_t = torch.empty(mstop, dtype=torch.int32, device=x.device)
        risk_t = torch.empty(mstop, dtype=torch.float64, device=x.device)
        P = _native.ptr
        rc = _native.cuda().cml_glmboost(P(xct), p, n, n_pad, P(ssd), P(rp), P(coef), int(mstop), float(nu), P(best), P(path_t), P(risk_t),
                                         _native.stream_ptr())
        _native.check(rc, "cml_glmboost")
        intercept = float(offset - (coef * mean).sum())
        return GLMBoostFit(intercept, coef, path_t.tolist(), risk_t.tolist())
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
