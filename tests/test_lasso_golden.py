"""Golden tests of the glmnet-semantics path solver against the reference's stored glmnet objects
(lasso_resultslist.rda, extracted into consensusml_b200/datasets/target_aml/glmnet_oracle.npz)."""
import numpy as np
import pytest
import torch

from consensusml_b200.models import lasso
from consensusml_b200.ops import lasso as K


def _ref_beta(o, tag, p=1984, L=100):
    ref = np.zeros((p, L))
    ref[o[tag + "_rows"]] = o[tag + "_beta"]
    return ref.T


def test_train_path_matches_glmnet(aml, glmnet_oracle):
    tr, _ = aml.split()
    x, y = tr.design(torch.float64), tr.labels().double()
    pb = K.lognet_path(x, y, torch.ones(1, x.shape[0], dtype=torch.float64), glmnet_oracle["rep1_lambda"])
    beta = pb.dense(0).numpy()
    ref = _ref_beta(glmnet_oracle, "rep1_train")
    assert np.abs(beta - ref).max() < 1e-10
    assert int(pb.npasses[0]) == int(glmnet_oracle["rep1_train_npasses"][0]) == 999
    assert np.abs(pb.dev_ratio[0].numpy() - glmnet_oracle["rep1_train_dev_ratio"]).max() < 1e-10
    assert np.array_equal(pb.df()[0].numpy(), glmnet_oracle["rep1_train_df"])
    assert abs(float(pb.nulldev[0]) - 128.92537558) < 1e-6


def test_run_lasso_reproduces_reference_rep1(aml, glmnet_oracle):
    res = lasso.run_lasso(aml)
    o = glmnet_oracle
    assert abs(res.cv.lambda_min - 0.16297508) < 1e-8
    assert abs(res.cv.lambda_1se - 0.65793322) < 1e-8
    assert np.abs(res.cv.cvm - o["rep1_cvm"]).max() < 1e-12
    assert np.abs(res.cv.cvsd - o["rep1_cvsd"]).max() < 1e-12
    assert np.array_equal(res.cv.nzero, o["rep1_cv_nzero"])
    assert res.confusion.tolist() == [[21, 2], [1, 20]]          # JSmith_code/Results/RiskGroup_Confusion_Matrix.csv
    assert abs(res.test_error - 0.06818182) < 1e-8
    assert abs(res.test_auc - 0.9835) < 1e-3                     # BASELINE.md derived AUC
    assert abs(res.test_logloss - 0.2122) < 1e-3
    assert res.final_npasses == 1128
    ref = dict(zip(o["rep1_nonzero_names"].tolist(), o["rep1_nonzero_coef"].tolist()))
    assert set(res.nonzero_coef) == set(ref) and len(ref) == 13
    for g, c in ref.items():
        assert abs(res.nonzero_coef[g] - c) < 1e-10


@pytest.mark.slow
def test_exclusion_reps_match_reference(aml, glmnet_oracle):
    reps = lasso.exclusion_reps(aml, nreps=4)
    expect_nnz = [13, 15, 24, 40]
    expect_err = [0.068, 0.068, 0.114, 0.068]
    for r, nnz, err in zip(reps, expect_nnz, expect_err):
        assert len(r.nonzero_coef) == nnz
        assert abs(r.test_error - err) < 1e-3


def test_elastic_net_and_intercept_kkt():
    """KKT conditions of the elastic-net logistic objective hold at the solution (closed-form check)."""
    g = torch.Generator().manual_seed(0)
    n, p = 60, 40
    x = torch.randn(n, p, generator=g, dtype=torch.float64)
    beta_true = torch.zeros(p, dtype=torch.float64)
    beta_true[:4] = torch.tensor([1.5, -2.0, 1.0, 0.5])
    y = (torch.rand(n, generator=g, dtype=torch.float64) < torch.sigmoid(x @ beta_true + 0.3)).double()
    lam = np.array([0.2, 0.1, 0.05, 0.02])
    for alpha in (1.0, 0.5):
        pb = K.lognet_path(x, y, torch.ones(1, n, dtype=torch.float64), lam, alpha=alpha, intercept=True, thresh=1e-12)
        for li, l in enumerate(lam):
            b = pb.dense(0)[li]
            a0 = pb.a0[0, li]
            mu = torch.sigmoid(x @ b + a0)
            grad = x.t() @ (y - mu) / n - l * (1 - alpha) * b
            nzm = b != 0
            assert torch.allclose(grad[nzm], l * alpha * torch.sign(b[nzm]), atol=2e-5)
            assert (grad[~nzm].abs() <= l * alpha + 2e-5).all()
            assert abs(float((y - mu).mean())) < 2e-5          # intercept stationarity


def test_feature_mask_and_weights_equivalence():
    """Zero weight == dropping the sample; masked-out feature == dropping the column."""
    g = torch.Generator().manual_seed(1)
    n, p = 40, 30
    x = torch.randn(n, p, generator=g, dtype=torch.float64)
    y = (torch.rand(n, generator=g) < 0.5).double()
    lam = np.array([0.1, 0.05])
    w = torch.ones(1, n, dtype=torch.float64)
    w[0, :5] = 0
    mask = torch.ones(p, dtype=torch.bool)
    mask[[3, 7]] = False
    a = K.lognet_path(x, y, w, lam, feature_mask=mask).dense(0)
    keep = torch.tensor([j for j in range(p) if j not in (3, 7)])
    bsub = K.lognet_path(x[5:][:, keep], y[5:], torch.ones(1, n - 5, dtype=torch.float64), lam).dense(0)
    assert torch.allclose(a[:, keep], bsub, atol=1e-12)
    assert (a[:, [3, 7]] == 0).all()


def test_stability_selection_cpu_small():
    g = torch.Generator().manual_seed(2)
    n, p = 50, 25
    x = torch.randn(n, p, generator=g, dtype=torch.float64)
    y = (x[:, 0] - x[:, 1] + 0.5 * torch.randn(n, generator=g, dtype=torch.float64) > 0).double()
    out = lasso.stability_selection(x, y, np.array([0.15, 0.08]), n_boot=40, seed=3)
    f = out["max_freq"]
    assert f[0] > 0.9 and f[1] > 0.9
    assert f[2:].mean() < 0.5


def test_covariance_active_sweeps_are_the_same_iteration(aml):
    """Host solver with covariance updates inside the active-set sweeps (no length-n reduction per coordinate step - the
    formulation evaluated for the GPU kernel): same pass counts and coefficients as the statement-for-statement glmnet loop."""
    import torch
    from consensusml_b200 import _native
    from consensusml_b200.ops import lasso as K
    from consensusml_b200.ops.replicate import replicate_weights
    tr, _ = aml.split()
    x, y = tr.design(torch.float64), tr.labels().double()
    w = torch.cat([torch.ones(1, 93, dtype=torch.float64), replicate_weights(3, 93, seed=1).double()[:, :93]])
    lam = K.default_lambda_grid()
    host = _native.host()
    try:
        host.cml_host_lognet_set_active_mode(1)
        cov = K.lognet_path(x, y, w, lam, want_eta=False)
    finally:
        host.cml_host_lognet_set_active_mode(0)
    ref = K.lognet_path(x, y, w, lam, want_eta=False)
    assert cov.npasses.tolist() == ref.npasses.tolist()
    for b in range(4):
        assert (cov.dense(b) - ref.dense(b)).abs().max() < 1e-12
