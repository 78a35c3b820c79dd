"""Normalisation: oracle self-checks on CPU, CUDA kernels vs the float64 oracle on GPU."""
import numpy as np
import pytest
import torch

from consensusml_b200.io.synthetic import nb_counts
from consensusml_b200.oracle import limma as O
from consensusml_b200.stats import normalize as N


@pytest.fixture(scope="module")
def small_counts():
    c, g = nb_counts(6000, 40, seed=3, n_de=300)
    return c, g


def test_tmm_geometric_mean_and_invariance(small_counts):
    c, _ = small_counts
    keep = O.cpm_filter(c, 1.0, 5)
    f = O.tmm_factors(c[keep].astype(float), c.sum(0).astype(float))
    assert abs(np.exp(np.log(f).mean()) - 1) < 1e-12
    perm = np.random.default_rng(0).permutation(c.shape[1])
    ref = O.tmm_ref_column(c[keep].astype(float), c.sum(0).astype(float))
    f2 = O.tmm_factors(c[keep][:, perm].astype(float), c.sum(0)[perm].astype(float), ref_column=int(np.flatnonzero(perm == ref)[0]))
    assert np.allclose(f2, f[perm], rtol=1e-10)


def test_tmm_identical_samples_give_unit_factors():
    rng = np.random.default_rng(1)
    base = rng.poisson(50, 3000)
    c = np.stack([base, base * 2, base * 3], 1)
    f = O.tmm_factors(c.astype(float))
    assert np.allclose(f, 1.0, atol=1e-12)


def test_logcpm_formula():
    c = np.array([[0, 10], [5, 0], [100, 200]], float)
    lib = np.array([1e6, 2e6])
    out = O.log_cpm(c, lib, np.array([1.0, 1.0]), 1.0)
    pcs = lib / lib.mean()
    assert np.allclose(out, np.log2((c + pcs) / (lib + 2 * pcs) * 1e6))


def test_cpu_api_matches_oracle(small_counts):
    c, _ = small_counts
    res = N.tmm_normalize(torch.from_numpy(c), min_samples=5)
    keep = O.cpm_filter(c, 1.0, 5)
    assert np.array_equal(res.keep.numpy(), keep)
    assert np.allclose(res.norm_factors.numpy(), O.tmm_factors(c[keep].astype(float), c.sum(0).astype(float)))


def test_lowess_matches_statsmodels_like_properties():
    rng = np.random.default_rng(4)
    x = np.sort(rng.uniform(0, 10, 400))
    y = np.sin(x) + 0.1 * rng.standard_normal(400)
    y[50] += 5                                # outlier is down-weighted by the robustness iterations
    fit = O.lowess(x, y, f=0.1)
    assert np.abs(fit - np.sin(x)).max() < 0.15
    # exact recovery of a straight line
    assert np.allclose(O.lowess(x, 2 * x + 1, f=0.5), 2 * x + 1, atol=1e-8)


def test_bh_matches_definition():
    p = np.array([0.01, 0.04, 0.03, 0.005, 0.8])
    adj = O.bh_adjust(p)
    o = np.argsort(p)
    manual = np.minimum.accumulate((p[o] * 5 / np.arange(1, 6))[::-1])[::-1]
    assert np.allclose(adj[o], np.minimum(manual, 1))


@pytest.mark.gpu
def test_gpu_normalisation_matches_oracle(small_counts):
    c, _ = small_counts
    ct = torch.from_numpy(c).cuda()
    lib = N.lib_sizes(ct)
    assert np.array_equal(lib.cpu().numpy(), c.sum(0).astype(float))
    keep = N.cpm_filter(ct, 1.0, 5, lib=lib)
    keep_ref = O.cpm_filter(c, 1.0, 5)
    assert np.array_equal(keep.cpu().numpy(), keep_ref)
    sub = ct[keep].contiguous()
    f = N.tmm_factors(sub, lib)
    f_ref = O.tmm_factors(c[keep_ref].astype(float), c.sum(0).astype(float))
    assert np.abs(f.cpu().numpy() - f_ref).max() < 2e-4, np.abs(f.cpu().numpy() - f_ref).max()
    lc = N.log_cpm(sub, lib, f)
    lc_ref = O.log_cpm(c[keep_ref].astype(float), c.sum(0).astype(float), f.cpu().numpy())
    assert np.abs(lc.cpu().numpy() - lc_ref).max() < 1e-4
    lct = N.log_cpm(sub, lib, f, sample_major=True)
    assert torch.equal(lct.t().contiguous(), lc)


@pytest.mark.gpu
def test_gpu_tmm_with_ties_and_zeros():
    rng = np.random.default_rng(9)
    c = rng.poisson(3, (5000, 12)).astype(np.int32)      # small counts: many tied M/A values, many zeros
    c[:, 3] *= 2
    ct = torch.from_numpy(c).cuda()
    f = N.tmm_factors(ct).cpu().numpy()
    f_ref = O.tmm_factors(c.astype(float))
    assert np.abs(f - f_ref).max() < 5e-4, (f, f_ref)
