"""Per-gene variance and top-k selection (K12).

Replaces ``ldf.var(axis=0)`` + ``sorted(...)[:1000]`` (scripts/model_comp.py:62-79) and
``order(rowVars(...))[1:100]`` (VikasP/AML.R:157-158).  The reduction is a single Welford-style pass per
column; when the cohort is sample-sharded the per-rank (count, mean, M2) triples are merged with Chan's
formula (see parallel/sharded.py) before the top-k."""
from __future__ import annotations

import torch


def column_moments(x: torch.Tensor):
    """(count, mean, M2) per column of a samples x genes matrix, in float64 - the mergeable summary."""
    xd = x.double()
    n = xd.shape[0]
    mean = xd.mean(0)
    m2 = ((xd - mean) ** 2).sum(0)
    return n, mean, m2


def merge_moments(a, b):
    """Chan et al. pairwise merge of two (count, mean, M2) summaries."""
    na, ma, sa = a
    nb, mb, sb = b
    n = na + nb
    d = mb - ma
    return n, ma + d * (nb / n), sa + sb + d * d * (na * nb / n)


def column_variance(x: torch.Tensor) -> torch.Tensor:
    n, _, m2 = column_moments(x)
    return m2 / (n - 1)


def variance_topk(x: torch.Tensor, k: int, largest: bool = True) -> torch.Tensor:
    """Indices of the k columns with the largest (or smallest) sample variance, ordered like the reference's
    ``sorted(..., reverse=...)`` (stable, ties by column index)."""
    v = column_variance(x)
    k = min(k, v.numel())
    order = torch.argsort(v, descending=largest, stable=True)
    return order[:k]
