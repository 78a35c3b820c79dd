"""GPU numerics: the warp-per-replicate sm_100a path kernel vs the float64 host oracle and the
reference's stored glmnet objects."""
import numpy as np
import pytest
import torch

from consensusml_b200.models import lasso
from consensusml_b200.ops import lasso as K

pytestmark = pytest.mark.gpu


def test_native_library_loaded():
    from consensusml_b200 import _native
    lib = _native.cuda()
    assert lib.cml_version() >= 100
    with open("/proc/self/maps") as fh:
        assert "libcml_b200.so" in fh.read()


def test_train_path_vs_glmnet(aml, glmnet_oracle):
    tr, _ = aml.split()
    x = tr.design(torch.float32).cuda()
    y = tr.labels().cuda()
    pb = K.lognet_path(x, y, torch.ones(1, x.shape[0], device="cuda"), glmnet_oracle["rep1_lambda"])
    torch.cuda.synchronize()
    assert int(pb.jerr[0]) == 0
    beta = pb.dense(0).double().cpu().numpy()
    ref = np.zeros((1984, 100))
    ref[glmnet_oracle["rep1_train_rows"]] = glmnet_oracle["rep1_train_beta"]
    ref = ref.T
    assert np.abs(beta - ref).max() < 2e-3          # fp32 vs glmnet's float64, same update sequence
    li = 89                                          # lambda.min
    assert set(np.flatnonzero(beta[li])) == set(np.flatnonzero(ref[li]))
    assert abs(int(pb.npasses[0]) - 999) <= 60
    assert np.abs(pb.dev_ratio[0].double().cpu().numpy() - glmnet_oracle["rep1_train_dev_ratio"]).max() < 1e-3


def test_batched_replicates_vs_host_oracle():
    g = torch.Generator().manual_seed(0)
    n, p, B = 75, 300, 40
    x = torch.randn(n, p, generator=g)
    beta = torch.zeros(p)
    beta[:6] = torch.tensor([1.0, -1.0, 0.8, -0.8, 0.5, 0.5])
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ beta)).float()
    w = torch.ones(B, n)
    for b in range(1, B):
        w[b, torch.randperm(n, generator=g)[:7]] = 0          # fold-like
    w[B - 1] = torch.bincount(torch.randint(0, n, (n,), generator=g), minlength=n).float()   # bootstrap-like
    mask = torch.ones(B, p, dtype=torch.bool)
    mask[3, :3] = False
    lam = np.geomspace(0.3, 0.01, 20)
    for kw in (dict(), dict(alpha=0.5, intercept=True), dict(modified_newton=True)):
        host = K.lognet_path(x.double(), y.double(), w.double(), lam, feature_mask=mask, **kw)
        dev = K.lognet_path(x.cuda(), y.cuda(), w.cuda(), lam, feature_mask=mask.cuda(), **kw)
        torch.cuda.synchronize()
        assert (dev.jerr == 0).all()
        for b in range(B):
            hb, db = host.dense(b), dev.dense(b).double().cpu()
            assert (hb - db).abs().max() < 5e-3, (kw, b, float((hb - db).abs().max()))
        assert (host.a0 - dev.a0.double().cpu()).abs().max() < 5e-3
        assert (host.eta - dev.eta.double().cpu()).abs().max() < 2e-2
        assert (host.dev_ratio - dev.dev_ratio.double().cpu()).abs().max() < 2e-3
        assert (dev.dense(3)[:, :3] == 0).all()


def test_run_lasso_on_gpu_reproduces_reference(aml):
    res = lasso.run_lasso(aml, device=torch.device("cuda"))
    assert abs(res.cv.lambda_min - 0.16297508) < 1e-6
    assert res.confusion.tolist() == [[21, 2], [1, 20]]
    assert abs(res.test_auc - 0.9835) < 0.01
    assert len(res.nonzero_coef) == 13


def test_all_padded_sizes_run():
    g = torch.Generator().manual_seed(5)
    lam = np.geomspace(0.3, 0.03, 8)
    for n in (20, 40, 90, 120, 150, 180, 250, 300, 500, 700, 1000):
        p = 64
        x = torch.randn(n, p, generator=g)
        y = (x[:, 0] + 0.5 * torch.randn(n, generator=g) > 0).float()
        host = K.lognet_path(x.double(), y.double(), torch.ones(2, n, dtype=torch.float64), lam)
        dev = K.lognet_path(x.cuda(), y.cuda(), torch.ones(2, n, device="cuda"), lam)
        torch.cuda.synchronize()
        assert (dev.jerr == 0).all()
        assert (host.dense(1) - dev.dense(1).double().cpu()).abs().max() < 5e-3, n


def test_stability_selection_freq_matches_dense():
    g = torch.Generator().manual_seed(7)
    n, p = 60, 100
    x = torch.randn(n, p, generator=g).cuda()
    y = (x[:, 0] - x[:, 1] > 0).float()
    lam = np.array([0.2, 0.1, 0.05])
    w = torch.ones(64, n, device="cuda")
    freq = torch.zeros(3, p, device="cuda")
    pb = K.lognet_path(x, y, w, lam, freq_out=freq)
    torch.cuda.synchronize()
    dense = torch.stack([pb.dense(b) for b in range(64)])
    assert torch.equal(freq, (dense != 0).sum(0).float())
