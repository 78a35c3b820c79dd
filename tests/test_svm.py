"""SVM: NumPy SMO restatement vs scikit-learn's libsvm (CPU), CUDA SMO + tcgen05 kernel matrix vs both (GPU),
the reference's stored linear-SVM weight column as a real-data oracle."""
import numpy as np
import pytest
import torch

from consensusml_b200.models import svm as S


def _xy(aml, device="cpu"):
    tr, te = aml.split()
    return (tr.design(torch.float32).to(device), tr.labels().to(device), te.design(torch.float32).to(device),
            te.labels().to(device))


def _sk(x, y, kernel):
    from sklearn.svm import SVC as SK
    xs = (x - x.mean(0)) / x.std(0, ddof=1)
    m = SK(kernel=kernel, C=1.0, gamma=1.0 / x.shape[1], tol=1e-3).fit(xs, y)
    return m, xs


@pytest.mark.parametrize("kernel,sk_kernel", [("linear", "linear"), ("radial", "rbf")])
def test_cpu_smo_matches_libsvm(aml, kernel, sk_kernel):
    xtr, ytr, xte, yte = _xy(aml)
    ours = S.SVC(kernel=kernel).fit(xtr, ytr)
    sk, xs = _sk(xtr.numpy().astype(np.float64), ytr.numpy(), sk_kernel)
    xte_s = (xte.numpy() - xtr.numpy().mean(0)) / xtr.numpy().std(0, ddof=1)
    d_ref = sk.decision_function(xte_s)
    d = ours.decision_function(xte).numpy()
    assert np.abs(d - d_ref).max() < 5e-3 * max(1.0, np.abs(d_ref).max())
    assert set(ours.support_.tolist()) == set(sk.support_.tolist())


def test_linear_weights_match_reference_column(aml):
    """standouttable.csv's svm1_weights column is t(coefs) %*% SV of the linear model (SEA:155,679-690)."""
    import os
    import pandas as pd
    path = "/root/reference/composite_code/rnotebook/data/standouttable.csv"
    if not os.path.exists(path):
        pytest.skip("reference tree not mounted")
    ref = pd.read_csv(path)
    xtr, ytr, *_ = _xy(aml)
    w = S.SVC(kernel="linear").fit(xtr, ytr).weights().numpy()
    rw = ref["svm1_weights"].to_numpy()
    c = np.corrcoef(w, rw)[0, 1]
    assert abs(c) > 0.999
    scale = np.abs(w).sum() / np.abs(rw).sum()
    assert abs(scale - 1) < 0.02                       # same magnitude; sign follows e1071's level order


def test_run_svm_protocol_and_tails(aml):
    xtr, ytr, xte, yte = _xy(aml)
    r1 = S.run_svm(xtr, ytr, xte, yte, "linear")
    r2 = S.run_svm(xtr, ytr, xte, yte, "linear", weightfilt=0.5)
    assert r1.columns.numel() == 1984 and r2.columns.numel() == 992
    assert torch.equal(r1.weightsvect, r2.weightsvect)          # the reference keeps the pre-filter weights (SEA:155)
    assert r1.test_error < 0.2 and r1.auc > 0.9
    tails = S.weight_quantile_tails(r1.weightsvect, aml.gene_ids)
    sizes = [len(tails["high"][f]) for f in (0.25, 0.10, 0.05, 0.01, 0.001)]
    assert sizes == [496, 199, 100, 20, 2]                      # 1984 genes: the reference's 1910-gene run gave 477/191/96/20/2


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["linear", "radial"])
def test_gpu_svm_matches_cpu(aml, kernel):
    xtr, ytr, xte, yte = _xy(aml)
    cpu = S.SVC(kernel=kernel).fit(xtr, ytr)
    gpu = S.SVC(kernel=kernel).fit(xtr.cuda(), ytr.cuda())
    dc = cpu.decision_function(xte)
    dg = gpu.decision_function(xte.cuda()).cpu()
    assert (dc - dg).abs().max() < 3e-2 * max(1.0, float(dc.abs().max()))       # bf16 kernel matrix
    assert ((dc > 0) == (dg > 0)).float().mean() > 0.97
    wc, wg = cpu.weights(), gpu.weights().cpu()
    assert float(torch.corrcoef(torch.stack([wc, wg]))[0, 1]) > 0.999


@pytest.mark.gpu
def test_gpu_grid_search_runs(aml):
    xtr, ytr, *_ = _xy(aml, "cuda")
    out = S.grid_search(xtr, ytr, costs=[0.1, 1.0, 10.0], gammas=[1e-4, 1.0 / 1984, 5e-3], nfolds=5)
    assert out["cv_error"].shape == (3, 3)
    assert float(out["cv_error"].min()) < 0.25


def _svm_blobs(n=420, p=12, seed=0):
    g = torch.Generator().manual_seed(seed)
    y = (torch.rand(n, generator=g) < 0.5).float()
    x = torch.randn(n, p, generator=g) + (y[:, None] - 0.5) * torch.tensor([1.6, -1.2, 0.8] + [0.0] * (p - 3))
    return x, y


@pytest.mark.parametrize("kernel", ["radial", "linear"])
def test_large_svm_decomposition_reaches_the_smo_optimum_cpu(kernel):
    """Working-set decomposition (models/svm_large.py) against the full-matrix SMO on a problem both can solve."""
    from consensusml_b200.models.svm_large import LargeSVC
    x, y = _svm_blobs()
    xs = (x - x.mean(0)) / x.std(0, unbiased=True)
    ref = S.SVC(kernel=kernel, scale=False).fit(xs, y)
    big = LargeSVC(kernel=kernel, q=64, tol=1e-3).fit(xs, y)
    ypm = 2 * y - 1
    K = ref._kernel(xs).double()
    def obj(a):
        ay = a.double() * ypm.double()
        return float(0.5 * ay @ K @ ay - a.double().sum())
    o_ref, o_big = obj((ref.coef_ * ypm).abs()), obj(big.alpha_)
    assert abs(o_big - o_ref) < 1e-3 * max(1.0, abs(o_ref)), (o_big, o_ref)
    assert abs(big.objective_ - o_big) < 1e-2 * max(1.0, abs(o_big))                  # the solver's own bookkeeping agrees
    assert big.gap_ < 1e-3 and 0 < big.outer_iters_ < 400
    d_ref, d_big = ref.decision_function(xs[:50]), big.decision_function(xs[:50])
    assert (d_ref - d_big).abs().max() < 5e-2 and ((d_ref > 0) == (d_big > 0)).float().mean() > 0.97
    assert big.cache_hit_rate_ > 0.3                                                    # working sets overlap: the LRU cache pays
