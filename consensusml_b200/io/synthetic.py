"""Synthetic cohorts of the reference's shapes.

The raw HTSeq count matrix (60,488 genes x 474 samples; JSmith_code/Differential_Expression_and_Lasso.Rmd:91,206)
is stripped from the reference tree, so count-level stages are exercised on negative-binomial counts of
the same shape, with a planted two-group signal so that the DE prefilter has something to find."""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np


def nb_counts(n_genes: int = 60488, n_samples: int = 145, *, frac_expressed: float = 0.32, n_de: int = 2000,
              lfc_sd: float = 1.6, dispersion: float = 0.25, seed: int = 0, frac_group1: float = 0.56,
              depth_range: Tuple[float, float] = (2e7, 8e7)) -> Tuple[np.ndarray, np.ndarray]:
    """Gene x sample integer counts and a 0/1 group label.

    ``frac_expressed`` of the genes carry appreciable expression (the reference keeps 18,243 of 60,488 genes
    after its CPM filter); ``n_de`` of those get a log2 fold change drawn from N(0, lfc_sd) between groups."""
    rng = np.random.default_rng(seed)
    group = (rng.random(n_samples) < frac_group1).astype(np.int8)
    expressed = rng.random(n_genes) < frac_expressed
    base = np.where(expressed, rng.lognormal(mean=-11.0, sigma=1.8, size=n_genes),
                    rng.lognormal(mean=-18.5, sigma=1.5, size=n_genes))
    base /= base.sum()
    lfc = np.zeros(n_genes)
    de = rng.choice(np.flatnonzero(expressed), size=min(n_de, int(expressed.sum())), replace=False)
    lfc[de] = rng.normal(0.0, lfc_sd, size=de.size)
    depth = rng.uniform(*depth_range, size=n_samples)
    mu = base[:, None] * (2.0 ** (lfc[:, None] * (group[None, :] - 0.5))) * depth[None, :]
    shape = 1.0 / dispersion
    lam = rng.gamma(shape, mu / shape)
    counts = rng.poisson(lam).astype(np.int32)
    return counts, group


def expression_cohort(n_samples: int, n_genes: int, n_signal: int = 100, seed: int = 0, effect=(0.2, 0.7),
                      frac_pos: float = 55 / 93):
    """log2-CPM-like matrix (samples x genes, fp32) with a weak diffuse class signal; mirrors the moments of
    the reference's modelling matrix (range about -7..15)."""
    rng = np.random.default_rng(seed)
    y = np.zeros(n_samples, np.float32)
    y[: int(round(n_samples * frac_pos))] = 1
    rng.shuffle(y)
    mu = rng.normal(2.0, 2.5, n_genes)
    sd = rng.uniform(0.6, 2.2, n_genes)
    x = (mu + sd * rng.standard_normal((n_samples, n_genes))).astype(np.float32)
    sig = rng.choice(n_genes, n_signal, replace=False)
    x[:, sig] += (y[:, None] - 0.5) * rng.uniform(*effect, n_signal).astype(np.float32) * sd[sig].astype(np.float32)
    return x, y


def nb_counts_multiclass(n_genes: int = 60488, class_sizes: Sequence[int] = (145, 160, 70, 130), *, frac_expressed: float = 0.35,
                         n_de_per_class: int = 1500, lfc_sd: float = 1.8, dispersion: float = 0.25, seed: int = 0,
                         depth_range: Tuple[float, float] = (2e7, 8e7)) -> Tuple[np.ndarray, np.ndarray]:
    """Pan-cancer-shaped counts: gene x sample integer matrix and an integer class label per sample.

    Defaults mirror the reference's pan-cancer matrix (60,488 genes x 474 TARGET samples across AML / NBL / WT / RT,
    GDC_Data_Download.Rmd:177-186; class sizes are placeholders of the right order, the real table is not on disk).
    Each class gets its own set of ``n_de_per_class`` marker genes shifted by N(0, lfc_sd) log2 units."""
    rng = np.random.default_rng(seed)
    label = np.repeat(np.arange(len(class_sizes)), class_sizes).astype(np.int64)
    rng.shuffle(label)
    n_samples = label.size
    expressed = rng.random(n_genes) < frac_expressed
    base = np.where(expressed, rng.lognormal(mean=-11.0, sigma=1.8, size=n_genes),
                    rng.lognormal(mean=-18.5, sigma=1.5, size=n_genes))
    base /= base.sum()
    log_mu = np.log2(base)[:, None] + np.zeros((1, n_samples))
    pool = np.flatnonzero(expressed)
    for c in range(len(class_sizes)):
        de = rng.choice(pool, size=min(n_de_per_class, pool.size), replace=False)
        shift = rng.normal(0.0, lfc_sd, size=de.size)
        log_mu[np.ix_(de, np.flatnonzero(label == c))] += shift[:, None]
    depth = rng.uniform(*depth_range, size=n_samples)
    mu = (2.0 ** log_mu) * depth[None, :]
    shape = 1.0 / dispersion
    counts = rng.poisson(rng.gamma(shape, mu / shape)).astype(np.int32)
    return counts, label
