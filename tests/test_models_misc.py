"""Discriminant classifiers, k-means, GLM, glmboost, MLP grid."""
import numpy as np
import pytest
import torch

from consensusml_b200.io.synthetic import nb_counts
from consensusml_b200.models import discriminant as D
from consensusml_b200.models import glm, glmboost, kmeans, mlp


@pytest.fixture(scope="module")
def counts100():
    c, g = nb_counts(3000, 96, seed=21, n_de=300, lfc_sd=2.0)
    c = torch.from_numpy(c).double() + 1                     # the reference adds a pseudo-count (VikasP/AML.R:167)
    var = torch.log2(c).var(1)
    top = torch.argsort(var, descending=True)[:100]          # top-100 variance genes (:157-158)
    return c[top], torch.from_numpy(g.astype(np.int64))


def test_size_factors_and_transforms(counts100):
    c, _ = counts100
    sf = D.deseq_size_factors(c)
    logc = np.log(c.numpy())
    ref = np.exp(np.median(logc - logc.mean(1, keepdims=True), axis=0))
    assert np.allclose(sf.numpy(), ref)
    doubled = c.clone()
    doubled[:, 0] *= 2
    assert abs(float(D.deseq_size_factors(doubled)[0] / sf[0]) - 2) < 0.2
    for f in (D.deseq_vst, D.deseq_rlog):
        t = f(c)
        assert torch.isfinite(t).all()
        hi = c.mean(1) > c.mean(1).median()
        # variance no longer grows with the mean: high- and low-expression genes have comparable spread
        assert 0.2 < float(t[hi].var(1).median() / t[~hi].var(1).median()) < 5


@pytest.mark.parametrize("method", ["PLDA", "PLDA2", "NBLDA", "pam", "voomNSC", "voomDLDA"])
def test_mlseq_methods_learn(counts100, method):
    c, y = counts100
    tr = torch.arange(96) % 10 < 7                             # 70 / 30 split (VikasP/AML.R:159-165)
    tune = {"PLDA": [0.0, 4.0], "PLDA2": [0.0, 4.0], "pam": [0.0, 1.5], "voomNSC": [0.0, 1.5]}.get(method)
    model, table = D.classify(c[:, tr], y[tr], method, tune=tune, folds=5, repeats=1)
    assert all(0.5 <= v <= 1.0 for v in table.values())
    cte = c[:, ~tr]
    if method in ("PLDA", "PLDA2", "NBLDA"):
        pred = model.predict(cte)
    elif method == "pam":
        t = D.deseq_vst(c, D.deseq_size_factors(c))
        pred = model.predict(t[:, ~tr])
    else:
        lib = cte.sum(0)
        pred = model.predict(torch.log2((cte + 0.5) / (lib + 1)[None, :] * 1e6))
    assert float((pred == y[~tr]).float().mean()) >= 0.8
    if hasattr(model, "selected_genes"):
        ids = [f"g{i}" for i in range(100)]
        assert 1 <= len(D.selected_genes(model, ids)) <= 100


def test_sparse_plda_selects_fewer_genes(counts100):
    c, y = counts100
    dense = D.PoissonLDA(rho=0.0).fit(c, y)
    sparse = D.PoissonLDA(rho=300.0).fit(c, y)
    assert sparse.selected_genes().numel() < dense.selected_genes().numel()


def test_kmeans_matches_sklearn_objective():
    from sklearn.cluster import KMeans
    rng = np.random.default_rng(1000)
    x = (rng.random((300, 35)) < rng.random((1, 35)) * 0.6).astype(np.float32)      # binary matrix like the reference's
    x[:100, :8] = 1
    x[100:200, 8:16] = 1
    ours = kmeans.kmeans(torch.from_numpy(x), 5, nstart=25)
    sk = KMeans(5, n_init=25, random_state=0).fit(x)
    assert ours.tot_withinss <= sk.inertia_ * 1.02
    assert abs(float(ours.withinss.sum()) - ours.tot_withinss) < 1e-2
    sweep = kmeans.kmeans_sweep(torch.from_numpy(x), ks=(6, 4, 2), nstart=5)
    assert sweep[2].tot_withinss > sweep[4].tot_withinss > sweep[6].tot_withinss
    # logistic GLM of a label on the cluster id (BuieRProj/KM_WHATEVER.Rmd:169-184)
    y = torch.from_numpy((np.arange(300) < 100).astype(np.float64))
    X = glm.design_from_factor(ours.labels, 5)
    fit = glm.logistic_glm(X, y)
    assert fit.deviance < fit.null_deviance * 0.5


def test_logistic_glm_matches_sklearn_unpenalised():
    from sklearn.linear_model import LogisticRegression
    g = torch.Generator().manual_seed(3)
    x = torch.randn(200, 4, generator=g, dtype=torch.float64)
    y = (torch.rand(200, generator=g, dtype=torch.float64) < torch.sigmoid(0.5 + x @ torch.tensor([1.0, -1.0, 0.5, 0.0], dtype=torch.float64))).double()
    X = torch.cat([torch.ones(200, 1, dtype=torch.float64), x], 1)
    fit = glm.logistic_glm(X, y)
    sk = LogisticRegression(C=1e10, tol=1e-10, max_iter=1000).fit(x.numpy(), y.numpy())
    assert np.allclose(fit.coef.numpy()[1:], sk.coef_[0], atol=1e-4) and abs(float(fit.coef[0]) - sk.intercept_[0]) < 1e-4
    assert fit.converged and (fit.se > 0).all()
    assert torch.allclose(glm.ensemble_mean([torch.tensor([0.2, 0.4]), torch.tensor([0.4, 0.8])]), torch.tensor([0.3, 0.6]))


def test_glmboost_recovers_signal_and_matches_definition():
    g = torch.Generator().manual_seed(4)
    x = torch.randn(132, 500, generator=g)
    y = 300 + 40 * x[:, 3] - 25 * x[:, 17] + 5 * torch.randn(132, generator=g)
    fit = glmboost.glmboost(x, y, mstop=100, nu=0.1)
    sel = set(fit.selected().tolist())
    assert {3, 17} <= sel and len(sel) < 60
    assert fit.risk[-1] < fit.risk[0] * 0.2 and all(b <= a + 1e-9 for a, b in zip(fit.risk, fit.risk[1:]))
    # first step by hand
    xc = x.double() - x.double().mean(0)
    r = y.double() - y.double().mean()
    j = int(((xc.t() @ r) ** 2 / (xc * xc).sum(0)).argmax())
    assert fit.path[0] == j
    m = glmboost.dichotomy_metrics(fit.predict(x), y, 300.0)
    assert m["tpr"] > 0.8 and m["mae"] < 30
    keep = glmboost.mean_difference_filter(x, y > 300, 0.5)
    assert 3 in keep.tolist()


def test_mlp_grid_cpu_trains_all_configs():
    g = torch.Generator().manual_seed(5)
    x = torch.randn(400, 20, generator=g)
    y = ((x[:, 0] - x[:, 1] + 0.5 * x[:, 2]) > 0).float()
    grid = mlp.make_grid(hidden=(8, 16), lr=(3e-3, 1e-2), weight_decay=(1e-3,))
    res = mlp.train_grid(x[:300], y[:300], grid, epochs=150, x_val=x[300:], y_val=y[300:])
    assert len(res.configs) == 4 and res.train_loss.shape == (4,)
    assert float(res.train_loss.min()) < 0.3 and float(res.val_auc.max()) > 0.9
    imp = res.input_importance(int(res.val_auc.argmax()))
    assert set(torch.argsort(-imp)[:3].tolist()) & {0, 1, 2}
    # the batched grid equals training one configuration alone
    solo = mlp.train_grid(x[:300], y[:300], [grid[2]], epochs=50)
    solo_seeded = mlp.train_grid(x[:300], y[:300], [dict(grid[2])], epochs=50)
    assert torch.allclose(solo.train_loss, solo_seeded.train_loss)


@pytest.mark.gpu
def test_mlp_grid_gpu_tcgen05(aml):
    tr, te = aml.split()
    xtr, ytr = tr.design(torch.float32).cuda(), tr.labels().cuda()
    xte, yte = te.design(torch.float32).cuda(), te.labels().cuda()
    grid = mlp.make_grid(hidden=(32, 64), lr=(1e-3, 3e-3), weight_decay=(0.0, 1e-3))
    res = mlp.train_grid(xtr, ytr, grid, epochs=120, x_val=xte, y_val=yte)
    assert float(res.val_auc.max()) > 0.9
    assert float(res.train_loss.max()) < 0.7


@pytest.mark.gpu
def test_kmeans_and_discriminant_on_gpu(counts100):
    c, y = counts100
    m_cpu = D.PoissonLDA(rho=2.0).fit(c, y)
    m_gpu = D.PoissonLDA(rho=2.0).fit(c.cuda(), y.cuda())
    assert torch.allclose(m_cpu.scores(c), m_gpu.scores(c.cuda()).cpu(), rtol=1e-9, atol=1e-6)
    x = torch.randn(400, 30)
    x[:200] += 2
    a, b = kmeans.kmeans(x, 2, nstart=5), kmeans.kmeans(x.cuda(), 2, nstart=5)
    assert abs(a.tot_withinss - b.tot_withinss) / a.tot_withinss < 1e-3


@pytest.mark.parametrize("method,tune", [("svmRadial", [0.25, 4.0]), ("LogitBoost", [11]), ("blackboost", [1]), ("deepboost", [2]),
                                          ("qda", [0.5])])
def test_mlseq_caret_learners_on_transformed_counts(counts100, method, tune):
    """svmRadial / LogitBoost / blackboost / deepboost trials of VikasP/AML.R:174-198,318-364 (+ the QDA stub of
    BuieRProj/KM_WHATEVER.Rmd:375-386): learners trained on the VST matrix, predicting from raw counts."""
    c, y = counts100
    tr = torch.arange(96) % 10 < 7
    model, table = D.classify(c[:, tr], y[tr], method, tune=tune, folds=2, repeats=1, pre="deseq-vst")
    assert all(0.5 <= v <= 1.0 for v in table.values()), table
    pred = model.predict(c[:, ~tr])
    assert float((pred == y[~tr]).float().mean()) >= 0.75
    assert 1 <= len(D.selected_genes(model, [f"g{i}" for i in range(100)])) <= 100


def test_sharded_solver_path_agrees_with_coordinate_descent():
    """FISTA path of the large-cohort solver (torch backend) against the glmnet-semantics path solver on the same grid:
    lambda_max gives the empty model, every later solution matches to optimisation tolerance."""
    from consensusml_b200.models import lasso_large as LL
    from consensusml_b200.ops import lasso as K
    g = torch.Generator().manual_seed(0)
    n, p = 400, 60
    x = torch.randn(n, p, generator=g)
    beta = torch.zeros(p)
    beta[:5] = torch.tensor([1.5, -1.2, 1.0, -0.8, 0.6])
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ beta)).float()
    lams, fits = LL.fit_path(x, y, n_lambda=6, max_iter=3000, tol=1e-7)
    assert int((fits[0].beta != 0).sum()) == 0 and int((fits[-1].beta != 0).sum()) > 5
    ref = K.lognet_path(x.double(), y.double(), torch.ones(1, n, dtype=torch.float64), lams, thresh=1e-12)
    for i in range(len(lams)):
        assert (fits[i].beta.double() - ref.dense(0)[i]).abs().max() < 5e-6
    with pytest.raises(ValueError):
        LL.fit_path(x, y, lams=[0.01, 0.1])
    # elastic net (glmnet's penalty, alpha = 0.5) against the same coordinate-descent solver
    lams, fits = LL.fit_path(x, y, n_lambda=5, max_iter=4000, tol=1e-7, alpha=0.5)
    ref = K.lognet_path(x.double(), y.double(), torch.ones(1, n, dtype=torch.float64), lams, alpha=0.5, thresh=1e-12)
    assert int((fits[0].beta != 0).sum()) == 0
    for i in range(len(lams)):
        assert (fits[i].beta.double() - ref.dense(0)[i]).abs().max() < 5e-6
