"""Unpenalised logistic regression by IRLS - R's ``glm(Risk ~ cluster, family = binomial)``
(BuieRProj/KM_WHATEVER.Rmd:169-184) - and the probability-averaging ensemble of scripts/model_comp.py:19-27."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence

import torch


@dataclass
class GLMFit:
    coef: torch.Tensor
    se: torch.Tensor
    z: torch.Tensor
    deviance: float
    null_deviance: float
    iters: int
    converged: bool


def design_from_factor(labels: torch.Tensor, n_levels: int) -> torch.Tensor:
    """Treatment contrasts: intercept + one indicator per level except the first."""
    oh = torch.nn.functional.one_hot(labels.long(), n_levels).double()
    return torch.cat([torch.ones(labels.numel(), 1, dtype=torch.float64, device=labels.device), oh[:, 1:]], 1)


def logistic_glm(x: torch.Tensor, y: torch.Tensor, max_iter: int = 25, tol: float = 1e-8) -> GLMFit:
    x = x.double()
    y = y.double()
    n, k = x.shape
    beta = torch.zeros(k, dtype=torch.float64, device=x.device)
    mu = (y + 0.5) / 2
    eta = torch.log(mu / (1 - mu))
    dev_old = float("inf")
    dev = dev_old
    cov = None
    it = 0
    ok = False
    for it in range(1, max_iter + 1):
        w = (mu * (1 - mu)).clamp_min(1e-10)
        zed = eta + (y - mu) / w
        xtw = x.t() * w
        cov = torch.linalg.pinv(xtw @ x)
        beta = cov @ (xtw @ zed)
        eta = x @ beta
        mu = torch.sigmoid(eta).clamp(1e-12, 1 - 1e-12)
        dev = float(-2 * (y * torch.log(mu) + (1 - y) * torch.log(1 - mu)).sum())
        if abs(dev - dev_old) / (abs(dev) + 0.1) < tol:
            ok = True
            break
        dev_old = dev
    p0 = y.mean().clamp(1e-12, 1 - 1e-12)
    null = float(-2 * (y * torch.log(p0) + (1 - y) * torch.log(1 - p0)).sum())
    se = torch.sqrt(torch.diagonal(cov))
    return GLMFit(beta, se, beta / se, dev, null, it, ok)


def ensemble_mean(probas: Sequence[torch.Tensor]) -> torch.Tensor:
    """Unweighted mean of class-1 probabilities (model_comp.py:19-27 averages RF, GB and XGB)."""
    return torch.stack([p.float() for p in probas]).mean(0)
