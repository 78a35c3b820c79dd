"""DESeq2 restatement (models/deseq.py, csrc/deseq.cu): gene-wise dispersions against a direct numerical maximisation of the
Cox-Reid adjusted profile likelihood, the Gamma-GLM trend against simulated truth, the rlog ridge fit against dense linear
algebra on the full [1 | I] design, the kernels (host execution model) against the vectorised float64 path."""
import numpy as np
import pytest
import torch
from scipy.optimize import minimize_scalar
from scipy.special import gammaln

from consensusml_b200 import _native
from consensusml_b200.models import deseq as D


def _simulate(G=240, S=30, seed=0, a0=0.05, a1=2.0):
    rng = np.random.default_rng(seed)
    mu = np.exp(rng.normal(4, 1.5, G))
    disp = a0 + a1 / mu
    sf = np.exp(rng.normal(0, 0.3, S))
    c = rng.negative_binomial(1 / disp[:, None], 1 / (1 + disp[:, None] * mu[:, None] * sf[None, :])).astype(np.float64)
    c[7] = 0                                                   # an all-zero gene
    return torch.from_numpy(c), disp


def _apl_numpy(log_alpha, y, mu, groups):
    al = np.exp(log_alpha)
    ll = np.sum(gammaln(y + 1 / al) - gammaln(1 / al) - y * np.log(mu + 1 / al) - np.log1p(mu * al) / al)
    w = 1 / (1 / mu + al)
    return ll - 0.5 * sum(np.log(w[groups == g].sum()) for g in np.unique(groups))


@pytest.mark.parametrize("n_groups", [1, 3])
def test_gene_dispersions_maximise_the_adjusted_profile_likelihood(n_groups):
    c, _ = _simulate(seed=n_groups)
    S = c.shape[1]
    groups = torch.from_numpy(np.arange(S) % n_groups)
    sf, _ = D.size_factors(c)
    bm, bv, disp, iters = D.gene_dispersions(c, sf, groups if n_groups > 1 else None)
    assert torch.isnan(disp[7]) and int(iters[7]) == 0 and float(bm[7]) == 0.0
    q = (c / sf[None, :]).numpy()
    assert np.allclose(bm.numpy(), q.mean(1)) and np.allclose(bv.numpy(), q.var(1, ddof=1))
    g = groups.numpy()
    checked = 0
    for i in range(0, 60, 3):
        if i == 7:
            continue
        mu = np.maximum(np.array([q[i][g == g[j]].mean() for j in range(S)]) * sf.numpy(), 0.5)
        res = minimize_scalar(lambda a: -_apl_numpy(a, c[i].numpy(), mu, g), bounds=(np.log(1e-8), np.log(max(10, S))), method="bounded",
                              options={"xatol": 1e-10})
        best = -res.fun
        ours = _apl_numpy(np.log(float(disp[i])), c[i].numpy(), mu, g)
        # the search stops at a likelihood change of 1e-6, and (DESeq2's rule) keeps its start unless the likelihood rose by a
        # millionth of its magnitude
        assert ours >= best - 2e-5 - abs(best) / 1e6, (i, ours, best)
        if 1e-6 < float(disp[i]) < 5 and abs(best) < 2000:
            assert abs(np.log(float(disp[i])) - res.x) < 0.05
        checked += 1
    assert checked >= 15


def test_parametric_trend_recovers_the_simulated_coefficients():
    c, disp_true = _simulate(G=1500, S=40, seed=5)
    fit = D.estimate_dispersions(c)
    assert fit.fit_type == "parametric"
    assert abs(fit.coefs[0] - 0.05) < 0.01 and abs(fit.coefs[1] - 2.0) < 0.3
    ok = torch.isfinite(fit.disp_fit)
    assert np.corrcoef(np.log(fit.disp_fit[ok].numpy()), np.log(disp_true[ok.numpy()]))[0, 1] > 0.97
    v, _ = D.vst(c, fit)
    q = c / fit.size_factors[None, :]
    big = q > 2000
    assert (v[big] - torch.log2(q[big])).abs().max() < 0.06    # the VST approaches log2 for large counts
    hi = fit.base_mean > 50
    assert float(v[hi].var(1).median()) < 2.0 * float(v[(fit.base_mean > 2) & (fit.base_mean < 20)].var(1).median()) + 0.5


def test_mean_and_local_trends_and_their_vst():
    c, _ = _simulate(G=400, S=24, seed=9, a0=0.2, a1=0.0)
    fm = D.estimate_dispersions(c, fit_type="mean")
    assert fm.fit_type == "mean" and abs(fm.coefs[0] - 0.2) < 0.06
    vm, _ = D.vst(c, fm)
    q = c / fm.size_factors[None, :]
    al = fm.coefs[0]
    assert torch.allclose(vm, (2 * torch.asinh(torch.sqrt(al * q)) - np.log(al) - np.log(4)) / np.log(2))
    fl = D.estimate_dispersions(c, fit_type="local")
    assert fl.fit_type == "local" and fl.trend is not None
    vl, _ = D.vst(c, fl)
    assert torch.isfinite(vl).all()
    order = torch.argsort(q[3])
    assert bool((vl[3][order][1:] >= vl[3][order][:-1] - 1e-9).all())     # monotone in the normalised count
    big = q > 3000
    if bool(big.any()):
        assert (vl[big] - torch.log2(q[big])).abs().max() < 0.5


def test_rlog_equals_the_dense_ridge_glm():
    c, _ = _simulate(G=60, S=12, seed=3)
    r = D.rlog(c)
    fit = r.dispersion
    S = c.shape[1]
    ln2 = np.log(2.0)
    lam = np.concatenate([[1e-6], np.full(S, 1.0 / r.beta_prior_var)]) / ln2 ** 2
    X = np.hstack([np.ones((S, 1)), np.eye(S)])
    sf = fit.size_factors.numpy()
    for i in (0, 3, 11, 25, 40):
        y, al = c[i].numpy(), float(fit.disp_fit[i])
        beta = np.concatenate([[np.log(float(fit.base_mean[i]))], np.zeros(S)])
        for _ in range(200):                                   # plain IRLS on the (S + 1)-column design, iterated to a fixed point
            mu = np.maximum(sf * np.exp(X @ beta), 0.5)
            w = mu / (1 + al * mu)
            z = np.log(mu / sf) + (y - mu) / mu
            new = np.linalg.solve(X.T @ (X * w[:, None]) + np.diag(lam), X.T @ (w * z))
            if np.abs(new - beta).max() < 1e-12:
                break
            beta = new
        assert np.abs((X @ beta) / ln2 - r.values[i].numpy()).max() < 2e-3      # ours stops at a relative deviance change of 1e-4
        assert abs(beta[0] / ln2 - float(r.intercept[i])) < 2e-3
    assert float(r.values[7].abs().max()) == 0.0 and np.isnan(float(r.intercept[7]))
    # shrinkage: rlog values are closer to the gene's mean than the plain log2 of the normalised counts, most of all at low counts
    q = c / fit.size_factors[None, :]
    plain = torch.log2(q + 0.5)
    low = (fit.base_mean > 0) & (fit.base_mean < 30)
    assert float(r.values[low].var(1).mean()) < float(plain[low].var(1).mean())


def test_weighted_quantile_matching():
    x = torch.randn(20000, generator=torch.Generator().manual_seed(1)) * 1.7
    v = D.match_weighted_upper_quantile_for_variance(x, torch.ones_like(x))
    assert abs(v - 1.7 ** 2) < 0.15
    w = torch.where(x.abs() > 3.0, torch.zeros_like(x), torch.ones_like(x))   # zero weight on the tails lowers the estimate
    assert D.match_weighted_upper_quantile_for_variance(x, w) < v


@pytest.mark.parametrize("n_groups", [1, 4])
def test_kernels_match_the_vectorised_path(n_groups):
    c, _ = _simulate(G=150, S=37, seed=20 + n_groups)
    groups = torch.from_numpy(np.arange(37) % n_groups) if n_groups > 1 else None
    ref = D.estimate_dispersions(c, groups=groups)
    rr = D.rlog(c, ref)
    fr = D.rlog(c[:, :5], frozen=rr)
    with _native.emulation(seed=3):
        _native.reset_calls()
        got = D.estimate_dispersions(c, groups=groups)
        gr = D.rlog(c, got)
        gf = D.rlog(c[:, :5], frozen=gr)
        assert _native.calls.get("cml_nb_dispersion") == 1 and _native.calls.get("cml_rlog") == 2
    ok = torch.isfinite(ref.disp_gene)
    assert torch.equal(torch.isfinite(got.disp_gene), ok)
    # same algorithm, different summation order: a gene whose last likelihood change sits at the 1e-6 threshold may take one
    # step more or less; everywhere else the iteration counts are identical and the estimates agree to rounding
    di = (got.iters.cpu() - ref.iters).abs()
    assert int(di.max()) <= 2 and float((di == 0).float().mean()) > 0.95
    rel = (got.disp_gene[ok] - ref.disp_gene[ok]).abs() / ref.disp_gene[ok]
    assert float(rel[di[ok] == 0].max()) < 1e-8 and float(rel.max()) < 2e-3
    assert got.fit_type == ref.fit_type and np.allclose(got.coefs, ref.coefs, rtol=1e-4)
    assert (gr.values - rr.values).abs().max() < 1e-3
    assert (gf.values - fr.values).abs().max() < 1e-3
    assert (fr.values - rr.values[:, :5]).abs().max() < 0.15   # frozen refit of training samples lands near their training values


@pytest.mark.parametrize("pre", ["deseq-vst", "deseq-rlog"])
def test_classifier_transforms_new_samples_with_the_training_fit(pre):
    """MLSeq's handling of test data: size factors against the training geometric means, the training dispersion function, and
    for rlog the training intercepts and prior variance - a single new sample can be transformed."""
    from consensusml_b200.models import discriminant as DC
    rng = np.random.default_rng(0)
    G, n = 100, 96
    y = torch.from_numpy((np.arange(n) % 2).astype(np.int64))
    mu = np.exp(rng.normal(4, 1, G))[:, None] * np.ones((1, n))
    mu[:15, y.numpy() == 1] *= 2.5
    c = torch.from_numpy(rng.negative_binomial(5, 5 / (5 + mu)).astype(np.float64) + 1)
    tr = torch.arange(n) % 10 < 7
    model, table = DC.classify(c[:, tr], y[tr], "svmRadial", tune=[1.0, 4.0], folds=5, repeats=1, pre=pre)
    pred = model.predict(c[:, ~tr])
    assert float((pred.cpu() == y[~tr]).float().mean()) >= 0.9
    one = model._transform(c[:, ~tr][:, :1])
    many = model._transform(c[:, ~tr])
    # a sample's transform does not depend on its batch (rlog: up to the IRLS stopping rule, which looks at the batch's deviance)
    assert one.shape == (1, G) and torch.allclose(one[0], many[0], atol=1e-4 if pre == "deseq-vst" else 0.05)
