"""Rank aggregation of per-model gene importances (north-star extension; BASELINE.json "consensus rank").

The reference stops at set intersections and ad-hoc thresholds (cml_targetaml_seanalysis.Rmd:1133-1141).
Here every model contributes a ranking of the genes by |importance| and the rankings are combined by
  * mean rank / Borda count,
  * robust rank aggregation (Kolde et al. 2012): for each gene the normalised ranks r_(1) <= ... <= r_(m) are
    compared with uniform order statistics, rho = min_k P(Beta(k, m-k+1) <= r_(k)), Bonferroni-corrected by m.
All computations are tensor ops on the device the importances live on (p x m is tiny next to the fits)."""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np
import pandas as pd
import torch
from scipy import stats

from ..ops.metrics import average_ranks


def rank_matrix(importance: torch.Tensor, absolute: bool = True, ties_zero_last: bool = True) -> torch.Tensor:
    """[p, m] importances -> [p, m] ranks (1 = most important, ties averaged).  Genes a model gives zero
    importance share the last (averaged) rank of that model."""
    imp = importance.abs() if absolute else importance
    cols = [average_ranks(-imp[:, j].double()) for j in range(imp.shape[1])]
    return torch.stack(cols, 1)


def mean_rank(ranks: torch.Tensor, weights: Optional[torch.Tensor] = None) -> torch.Tensor:
    if weights is None:
        return ranks.mean(1)
    w = weights.double() / weights.sum()
    return (ranks * w[None, :]).sum(1)


def borda(ranks: torch.Tensor) -> torch.Tensor:
    """Borda score: sum over models of (p - rank); larger is better."""
    return (ranks.shape[0] - ranks).sum(1)


def robust_rank_aggregation(ranks: torch.Tensor) -> torch.Tensor:
    """RRA rho score per gene (smaller = more consistently top-ranked), in [0, 1]."""
    p, m = ranks.shape
    r = torch.sort(ranks.double() / p, dim=1).values.cpu().numpy()
    k = np.arange(1, m + 1)
    pv = stats.beta.cdf(r, k[None, :], (m - k + 1)[None, :])
    rho = np.minimum(pv.min(1) * m, 1.0)
    return torch.as_tensor(rho, device=ranks.device)


def aggregate(importances: Dict[str, Sequence[float]], gene_ids: Sequence[str], method: str = "rra",
              top: Optional[int] = None) -> pd.DataFrame:
    """Consensus table: one row per gene, per-model ranks, mean rank, Borda, RRA rho; sorted by ``method``."""
    names = list(importances)
    imp = torch.stack([torch.as_tensor(np.array(importances[n], dtype=np.float64)) for n in names], 1)
    ranks = rank_matrix(imp)
    df = pd.DataFrame({f"rank_{n}": ranks[:, j].numpy() for j, n in enumerate(names)}, index=list(gene_ids))
    df["mean_rank"] = mean_rank(ranks).numpy()
    df["borda"] = borda(ranks).numpy()
    df["rra_rho"] = robust_rank_aggregation(ranks).numpy()
    key = {"rra": ("rra_rho", True), "mean_rank": ("mean_rank", True), "borda": ("borda", False)}[method]
    df = df.sort_values(key[0], ascending=key[1], kind="stable")
    return df.head(top) if top else df


def top_k_overlap(a: Sequence[str], b: Sequence[str], k: int = 50) -> float:
    """|top-k(a) & top-k(b)| / k - the BASELINE.json "top-50 gene overlap" acceptance metric."""
    return len(set(list(a)[:k]) & set(list(b)[:k])) / float(k)
