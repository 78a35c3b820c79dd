"""Evaluation metrics (K18): confusion matrix, error, TPR/TNR/FDR/FOR, log-loss, ROC/AUC, chi-square.

Reference: confusion + test error cml_targetaml_seanalysis.Rmd:105-108; TPR/TNR/FDR/FOR :815-848;
log-loss scripts/model_comp.py:24-27; ROC via ROCR :176-177; chi-square balance checks :462,470.
The reference never prints an AUC; BASELINE.md derives 0.9835 from the stored model, so AUC is
first-class here.  These are tiny reductions (n <= a few hundred in the reference, counts are
all-reduced across ranks when sample-sharded), implemented with torch tensor ops on whatever
device the inputs live on.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict

import torch


@dataclass
class BinaryReport:
    confusion: torch.Tensor   # [[TN, FN], [FP, TP]] laid out like R's table(pred, truth)
    error: float
    tpr: float
    tnr: float
    fdr: float
    fomr: float               # false omission rate ("FOR" in the reference)
    precision: float
    recall: float

    def as_dict(self) -> Dict[str, float]:
        return {"error": self.error, "tpr": self.tpr, "tnr": self.tnr, "fdr": self.fdr, "for": self.fomr,
                "precision": self.precision, "recall": self.recall}


def confusion_matrix(pred: torch.Tensor, truth: torch.Tensor) -> torch.Tensor:
    """2x2 table with rows = predicted class, columns = true class (R's ``table(pred, truth)``)."""
    pred = pred.long().flatten()
    truth = truth.long().flatten()
    idx = pred * 2 + truth
    return torch.bincount(idx, minlength=4).reshape(2, 2)


def _ratio(a: float, b: float) -> float:
    return a / b if b > 0 else float("nan")


def binary_report(pred: torch.Tensor, truth: torch.Tensor) -> BinaryReport:
    cm = confusion_matrix(pred, truth)
    tn, fn, fp, tp = (float(v) for v in cm.flatten().tolist())
    n = tn + fn + fp + tp
    return BinaryReport(cm, _ratio(fn + fp, n), _ratio(tp, tp + fn), _ratio(tn, tn + fp), _ratio(fp, fp + tp),
                        _ratio(fn, fn + tn), _ratio(tp, tp + fp), _ratio(tp, tp + fn))


def log_loss(truth: torch.Tensor, prob: torch.Tensor, eps: float = 1e-15) -> float:
    """Binary cross-entropy of P(class 1), clipped like scikit-learn's ``log_loss``."""
    if prob.dim() == 2:
        prob = prob[:, 1]
    p = prob.double().clamp(eps, 1 - eps)
    t = truth.double()
    return float(-(t * p.log() + (1 - t) * (1 - p).log()).mean())


def roc_curve(score: torch.Tensor, truth: torch.Tensor):
    """(fpr, tpr, thresholds) at every distinct score, descending (ROCR ``performance(pred,'tpr','fpr')``)."""
    s = score.double().flatten()
    t = truth.double().flatten()
    order = torch.argsort(s, descending=True, stable=True)
    s, t = s[order], t[order]
    tp = torch.cumsum(t, 0)
    fp = torch.cumsum(1 - t, 0)
    distinct = torch.ones_like(s, dtype=torch.bool)
    distinct[:-1] = s[1:] != s[:-1]
    tp, fp, thr = tp[distinct], fp[distinct], s[distinct]
    P, N = t.sum(), (1 - t).sum()
    zero = torch.zeros(1, dtype=torch.float64, device=s.device)
    tpr = torch.cat([zero, tp / P])
    fpr = torch.cat([zero, fp / N])
    thr = torch.cat([torch.full((1,), float("inf"), dtype=torch.float64, device=s.device), thr])
    return fpr, tpr, thr


def auc(score: torch.Tensor, truth: torch.Tensor) -> float:
    """Area under the ROC curve by the rank-sum (Mann-Whitney) identity with average ranks for ties."""
    s = score.double().flatten()
    t = truth.double().flatten()
    n1 = float(t.sum())
    n0 = float((1 - t).sum())
    if n1 == 0 or n0 == 0:
        return float("nan")
    r = average_ranks(s)
    return float((r[t > 0.5].sum() - n1 * (n1 + 1) / 2) / (n1 * n0))


def average_ranks(v: torch.Tensor) -> torch.Tensor:
    """1-based ranks with ties sharing their average (R's default ``rank``)."""
    v = v.flatten()
    order = torch.argsort(v, stable=True)
    sv = v[order]
    n = v.numel()
    pos = torch.arange(1, n + 1, dtype=torch.float64, device=v.device)
    new_group = torch.ones(n, dtype=torch.bool, device=v.device)
    new_group[1:] = sv[1:] != sv[:-1]
    gid = torch.cumsum(new_group.long(), 0) - 1
    ng = int(gid[-1].item()) + 1 if n else 0
    gsum = torch.zeros(ng, dtype=torch.float64, device=v.device).index_add_(0, gid, pos)
    gcnt = torch.zeros(ng, dtype=torch.float64, device=v.device).index_add_(0, gid, torch.ones_like(pos))
    ranks_sorted = (gsum / gcnt)[gid]
    out = torch.empty(n, dtype=torch.float64, device=v.device)
    out[order] = ranks_sorted
    return out


def chisq_test(table: torch.Tensor, correct: bool = True):
    """Pearson chi-square test of independence on an r x c table (R ``chisq.test``; Yates for 2x2).
    Returns (statistic, dof, p-value)."""
    from scipy.stats import chi2
    o = table.double()
    n = o.sum()
    e = o.sum(1, keepdim=True) * o.sum(0, keepdim=True) / n
    if correct and tuple(o.shape) == (2, 2):
        yates = torch.minimum(torch.full_like(o, 0.5), (o - e).abs())
        stat = (((o - e).abs() - yates) ** 2 / e).sum()
    else:
        stat = ((o - e) ** 2 / e).sum()
    dof = (o.shape[0] - 1) * (o.shape[1] - 1)
    return float(stat), dof, float(chi2.sf(float(stat), dof))
