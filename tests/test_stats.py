import numpy as np
import pytest
import torch
from scipy import stats as sps

from consensusml_b200.stats import correlation as C


def test_cpu_spearman_matches_scipy(aml):
    x = aml.assay[:60].float()
    ours = C.correlation(x).numpy()
    ref = sps.spearmanr(x.numpy().T).correlation
    assert np.abs(ours - ref).max() < 1e-6
    sub = list(range(0, 137, 3))
    ours = C.correlation(x, samples=sub, method="pearson").numpy()
    assert np.abs(ours - np.corrcoef(x.numpy()[:, sub])).max() < 1e-6


def test_cpu_kde_pca_hclust():
    g = torch.Generator().manual_seed(0)
    v = torch.randn(500, generator=g)
    grid, d = C.kde(v)
    ref = sps.gaussian_kde(v.numpy(), bw_method=C.bw_nrd0(v) / v.numpy().std(ddof=1))(grid.numpy())
    assert np.abs(d.numpy() - ref).max() < 1e-4
    assert abs(float(torch.trapezoid(d, grid)) - 1) < 1e-3
    x = torch.randn(40, 300, generator=g)
    x[:20, :50] += 3
    sc, evr = C.pca2(x)
    u, s, _ = np.linalg.svd((x - x.mean(0)).numpy(), full_matrices=False)
    assert np.allclose(np.abs(sc.numpy()), np.abs(u[:, :2] * s[:2]), atol=1e-3)
    assert (sc[:20, 0].mean() - sc[20:, 0].mean()).abs() > 5
    ro, co = C.heatmap_order(x[:, :60])
    assert sorted(ro.tolist()) == list(range(40))
    blocks = np.array([r < 20 for r in ro])
    assert (np.diff(blocks.astype(int)) != 0).sum() <= 1          # the two planted groups come out contiguous


@pytest.mark.gpu
def test_gpu_spearman_and_kde(aml):
    x = aml.assay.float().cuda()
    ex = C.correlation(x[:400], exact=True).cpu().numpy()
    ref = sps.spearmanr(aml.assay[:400].float().numpy().T).correlation      # same fp32 values, hence the same ties
    assert np.abs(ex - ref).max() < 2e-5
    tc = C.correlation(x).cpu()                                    # 1984 x 1984 through the tcgen05 GEMM
    exf = C.correlation(x, exact=True).cpu()
    assert (tc - exf).abs().max() < 8e-3
    assert (tc.diagonal() - 1).abs().max() < 8e-3
    sub = list(range(0, 137, 2))
    a = C.correlation(x[:200], x[200:500], samples=sub, exact=True).cpu().numpy()
    full = sps.spearmanr(aml.assay[:500].float().numpy()[:, sub].T).correlation
    assert np.abs(a - full[:200, 200:500]).max() < 2e-5
    v = torch.randn(20000, device="cuda")
    g1, d1 = C.kde(v)
    g2, d2 = C.kde(v.cpu())
    assert (d1.cpu() - d2).abs().max() < 1e-4


@pytest.mark.parametrize("method", ["complete", "average", "single", "weighted", "ward"])
def test_own_hclust_matches_scipy_linkage(method):
    """stats.correlation.hclust is the framework's nearest-neighbour-chain implementation; scipy is only the oracle."""
    from scipy.cluster.hierarchy import leaves_list, linkage
    from scipy.spatial.distance import pdist, squareform
    rng = np.random.default_rng(3)
    for n in (2, 3, 17, 120):
        x = rng.standard_normal((n, 6))
        d = squareform(pdist(x))
        Z, leaves = C.hclust(torch.from_numpy(d), method)
        ref = linkage(pdist(x), method=method)
        assert np.allclose(Z[:, 2], ref[:, 2], rtol=1e-10, atol=1e-12), method
        assert np.array_equal(Z[:, [0, 1, 3]], ref[:, [0, 1, 3]]), method
        assert np.array_equal(leaves, leaves_list(ref))
    assert "scipy" not in C.hclust.__code__.co_names and "linkage" not in C.hclust_order.__code__.co_names
