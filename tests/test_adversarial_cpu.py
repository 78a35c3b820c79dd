"""Degenerate inputs the real data never shows but a library must survive: constant or duplicated genes, all-zero rows and
columns with massive ties, a class with a single member.  CPU paths (the oracles the GPU kernels are compared with)."""
import numpy as np
import pytest
import torch

from consensusml_b200.ops import lasso as K
from consensusml_b200.oracle import limma as O


def test_lasso_ignores_constant_genes_and_splits_duplicates_deterministically():
    g = torch.Generator().manual_seed(1)
    n, p = 60, 12
    x = torch.randn(n, p, generator=g, dtype=torch.float64)
    x[:, 3] = 2.5                                  # constant gene: can never enter (glmnet's chkvars)
    x[:, 7] = x[:, 0]                              # exact duplicate of an informative gene
    y = (x[:, 0] + 0.3 * torch.randn(n, generator=g, dtype=torch.float64) > 0).double()
    lam = 10.0 ** np.linspace(0, -2.5, 20)
    out = K.lognet_path(x, y, torch.ones(1, n, dtype=torch.float64), lam)
    beta = out.dense(0)
    assert int(out.jerr[0]) == 0 and torch.isfinite(beta).all()
    assert (beta[:, 3] == 0).all()
    # cyclic order: the first of two identical columns takes (essentially all of) the coefficient
    assert (beta[-1, 0] != 0) and beta[:, 7].abs().max() < 0.05 * beta[:, 0].abs().max()
    w = torch.ones(1, n, dtype=torch.float64)
    w[0, 10:] = 0                                                   # 10 samples left, all of one class after masking?
    y2 = y.clone()
    y2[:10] = 1.0
    bad = K.lognet_path(x, y2, w, lam)
    assert int(bad.jerr[0]) in (8001, 9001)                         # glmnet's "one class has (almost) no weight" codes


def test_tmm_and_cpm_with_zero_genes_zero_heavy_samples_and_ties():
    rng = np.random.default_rng(2)
    c = rng.poisson(20, size=(400, 8)).astype(np.float64)
    c[:50] = 0                                                      # genes with no reads anywhere
    c[50:120, 3] = 0                                                # a sample with many drop-outs
    c[120:200] = 7                                                  # massive ties
    f = O.tmm_factors(c)
    assert np.isfinite(f).all() and abs(np.exp(np.log(f).mean()) - 1) < 1e-12
    lc = O.log_cpm(c, c.sum(0) * f, prior_count=1.0)
    assert np.isfinite(lc).all()
    keep = O.cpm_filter(c, min_cpm=1.0, min_samples=2)
    assert not keep[:50].any() and keep[120:200].all()


def test_voom_de_survives_a_constant_gene_and_tiny_groups():
    from consensusml_b200.io.synthetic import nb_counts
    c, g = nb_counts(800, 14, seed=4, n_de=80)
    c = c.astype(np.float64)
    c[5] = 1000                                                     # constant, highly expressed gene: zero residual variance
    g[:] = 0
    g[:2] = 1                                                       # a group of two
    r = O.voom_de(c, g, 0.05)
    eb = r["ebayes"]
    assert np.isfinite(eb.t[np.isfinite(eb.t)]).all() and np.isfinite(eb.p_value).sum() > 0.9 * eb.p_value.size
    adj = O.bh_adjust(np.nan_to_num(eb.p_value, nan=1.0))
    assert ((adj >= 0) & (adj <= 1)).all()


def test_binner_rank_correlation_and_kmeans_on_ties():
    from consensusml_b200.models import kmeans
    from consensusml_b200.models.trees import Binner
    from consensusml_b200.stats import correlation as C
    x = torch.zeros(30, 5)
    x[:, 1] = torch.arange(30).float()
    x[:, 2] = (torch.arange(30) % 2).float()
    x[:, 3] = torch.arange(30).float().flip(0)
    b, bins = Binner.fit(x)
    assert int(bins[0].max()) == 0 and int(bins[2].max()) == 1      # constant feature: one bin; binary feature: two
    r = C.correlation(x[:, 1:4].t().contiguous()).numpy()           # rows = genes
    assert abs(r[0, 2] + 1) < 1e-6 and abs(r[0, 0] - 1) < 1e-6      # perfectly anti-correlated ranks; ties averaged
    assert np.isfinite(r).all()
    pts = torch.cat([torch.zeros(10, 2), torch.ones(10, 2) * 5])     # two point masses
    res = kmeans.kmeans(pts, 2, nstart=3)
    assert float(res.tot_withinss) < 1e-9 and len(set(res.labels.tolist())) == 2
