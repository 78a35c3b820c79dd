"""Evaluation metrics (K18): confusion matrix, error, TPR/TNR/FDR/FOR, log-loss, ROC/AUC, chi-square.

Reference: confusion + test error cml_targetaml_seanalysis.Rmd:105-108; TPR/TNR/FDR/FOR :815-848;
log-loss scripts/model_comp.py:24-27; ROC via ROCR :176-177; chi-square balance checks :462,470.
The reference never prints an AUC; BASELINE.md derives 0.9835 from the stored model, so AUC is
first-class here.  ``batch_metrics`` evaluates many score vectors (every lambda of a path, every grid
configuration, every replicate) in ONE launch of ``binary_metrics_kernel`` (csrc/metrics.cu: in-CTA sort,
rank-sum AUC with tie handling, log-loss, confusion counts); ``auc`` / ``log_loss`` / ``confusion_matrix`` of a
device tensor go through it too.  CPU tensors use the torch restatement below, which is also the oracle.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Dict, Optional

import torch

from .. import _native

_native.register_signature("cml_binary_metrics", [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                                  ctypes.c_longlong, ctypes.c_float, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p])
_KERNEL_MAX_N = 8192


_KERNEL_STATE: dict = {}
_FORCE_CHECK = False            # tests: run the first-use comparison on the host execution model too


def _batch_metrics_torch(scores: torch.Tensor, truth: torch.Tensor, threshold: float, is_prob: bool) -> torch.Tensor:
    B = scores.shape[0]
    out = torch.empty(B, 8, dtype=torch.float64, device=scores.device)
    for b in range(B):
        t = truth if truth.dim() == 1 else truth[b]
        cm = _confusion_torch((scores[b] > threshold), t)
        out[b, 0] = _auc_torch(scores[b], t)
        out[b, 1] = _log_loss_torch(t, scores[b]) if is_prob else float("nan")
        out[b, 2:6] = cm.flatten().double()
        out[b, 6], out[b, 7] = cm[:, 1].sum(), cm[:, 0].sum()
    return out


def batch_metrics(scores: torch.Tensor, truth: torch.Tensor, threshold: float = 0.0, is_prob: bool = False) -> Dict[str, torch.Tensor]:
    """Metrics of B score vectors at once.  scores [B, n]; truth [n] or [B, n] in {0, 1}; class 1 iff score > threshold
    (0 for decision values / link scale, 0.5 for probabilities).  Returns float64 tensors of length B:
    auc, logloss (NaN unless ``is_prob``), tn, fn, fp, tp, error."""
    if scores.dim() == 1:
        scores = scores[None, :]
    B, n = scores.shape
    if _native.on_device(scores) and n <= _KERNEL_MAX_N and _KERNEL_STATE.get("use", True):
        sc = scores.float().contiguous()
        tr = truth.to(device=sc.device, dtype=torch.float32).contiguous()
        out = torch.empty(B, 8, dtype=torch.float64, device=sc.device)
        rc = _native.cuda().cml_binary_metrics(_native.ptr(sc), sc.stride(0), B, n, _native.ptr(tr), n if tr.dim() == 2 else 0,
                                               float(threshold), int(is_prob), _native.ptr(out), _native.stream_ptr())
        _native.check(rc, "cml_binary_metrics")
        if (scores.is_cuda or _FORCE_CHECK) and "checked" not in _KERNEL_STATE:
            # first use on a real device: this kernel has run on the host execution model only - hold its first answer against
            # the torch restatement, and fall back to that (loudly) if they disagree
            _KERNEL_STATE["checked"] = True
            ref = _batch_metrics_torch(scores, truth, threshold, is_prob)
            same = torch.allclose(out[:, 0], ref[:, 0], atol=1e-6, equal_nan=True) and torch.equal(out[:, 2:6], ref[:, 2:6]) and \
                torch.allclose(out[:, 1], ref[:, 1], rtol=1e-5, atol=1e-6, equal_nan=True)
            if not same:
                import warnings
                warnings.warn("consensusml_b200: cml_binary_metrics disagreed with the torch restatement on this device; using the latter")
                _KERNEL_STATE["use"] = False
                out = ref
    else:
        out = _batch_metrics_torch(scores, truth, threshold, is_prob)
    return {"auc": out[:, 0], "logloss": out[:, 1], "tn": out[:, 2], "fn": out[:, 3], "fp": out[:, 4], "tp": out[:, 5],
            "error": (out[:, 3] + out[:, 4]) / n}


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
    if _native.on_device(pred) and pred.numel() <= _KERNEL_MAX_N:
        m = batch_metrics(pred.flatten().float(), truth.flatten(), threshold=0.5)
        return torch.stack([m["tn"], m["fn"], m["fp"], m["tp"]], 1).reshape(2, 2).long()
    return _confusion_torch(pred, truth)


def _confusion_torch(pred: torch.Tensor, truth: torch.Tensor) -> torch.Tensor:
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
    if _native.on_device(prob) and prob.numel() <= _KERNEL_MAX_N and eps == 1e-15:
        return float(batch_metrics(prob.flatten(), truth.flatten(), threshold=0.5, is_prob=True)["logloss"][0])
    return _log_loss_torch(truth, prob, eps)


def _log_loss_torch(truth: torch.Tensor, prob: torch.Tensor, eps: float = 1e-15) -> float:
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
    if _native.on_device(score) and score.numel() <= _KERNEL_MAX_N:
        return float(batch_metrics(score.flatten(), truth.flatten())["auc"][0])
    return _auc_torch(score, truth)


def _auc_torch(score: torch.Tensor, truth: torch.Tensor) -> float:
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
