"""limma-voom DE: oracle self-consistency on CPU, CUDA kernels vs the oracle on GPU, DEG table format."""
import io

import numpy as np
import pytest
import torch

from consensusml_b200.io.synthetic import nb_counts
from consensusml_b200.oracle import limma as O
from consensusml_b200.stats import de as DE


@pytest.fixture(scope="module")
def cohort():
    c, g = nb_counts(8000, 60, seed=11, n_de=400)
    return c, g


def test_oracle_finds_planted_signal_and_moderation_is_sane(cohort):
    c, g = cohort
    r = O.voom_de(c, g)
    eb = r["ebayes"]
    assert 50 < len(r["deg_index"]) < 600
    assert eb.df_prior > 1 and 0.3 < eb.s2_prior < 3
    # moderated variances sit between the gene-wise estimate and the prior
    s2 = r["fit"].sigma ** 2
    lo, hi = np.minimum(s2, eb.s2_prior), np.maximum(s2, eb.s2_prior)
    assert np.all(eb.s2_post >= lo - 1e-12) and np.all(eb.s2_post <= hi + 1e-12)
    adj = O.bh_adjust(eb.p_value)
    assert np.all(adj >= eb.p_value - 1e-15)
    # WLS with unit weights == OLS
    e = r["E"]
    design = np.stack([g == 1, g == 0], 1).astype(float)
    a, b = O.lm_fit(e, design), O.lm_fit(e, design, np.ones_like(e))
    assert np.allclose(a.coefficients, b.coefficients) and np.allclose(a.sigma, b.sigma)
    # group-means design: coefficient = weighted group mean
    w = r["weights"]
    m1 = (w * e)[:, g == 1].sum(1) / w[:, g == 1].sum(1)
    assert np.allclose(r["fit"].coefficients[:, 0], m1)


def test_cpu_api_and_table_format(cohort, tmp_path):
    c, g = cohort
    res = DE.voom_de(torch.from_numpy(c), g)
    ids = [f"ENSG{i:011d}.1" for i in range(c.shape[0])]
    tab = res.table(ids)
    assert list(tab.columns) == ["logFC", "AveExpr", "t", "P.Value", "adj.P.Val", "B"]
    assert (tab["adj.P.Val"] < 0.05).all() and (tab["logFC"].abs() > 1).all()
    assert (np.diff(tab["P.Value"].to_numpy()) >= 0).all()
    path = tmp_path / "degs.csv"
    DE.write_deg_table(tab, str(path))
    first = path.read_text().splitlines()[0]
    assert first == '"","logFC","AveExpr","t","P.Value","adj.P.Val","B"'


@pytest.mark.gpu
def test_gpu_voom_de_matches_oracle(cohort):
    c, g = cohort
    ref = O.voom_de(c, g)
    res = DE.voom_de(torch.from_numpy(c).cuda(), g)
    assert np.array_equal(res.keep.cpu().numpy(), ref["keep"])
    assert np.abs(res.norm_factors.cpu().numpy() - ref["norm_factors"]).max() < 3e-4, 'norm factors'
    eb = ref["ebayes"]
    assert np.abs(res.logFC - ref["coef"]).max() < 2e-3, ('logFC', np.abs(res.logFC - ref['coef']).max())
    assert np.abs(res.AveExpr - ref["fit"].amean).max() < 1e-3, 'AveExpr'
    assert abs(res.s2_prior - eb.s2_prior) / eb.s2_prior < 5e-3, ('s2_prior', res.s2_prior, eb.s2_prior)
    assert abs(res.df_prior - eb.df_prior) / eb.df_prior < 2e-2, ('df_prior', res.df_prior, eb.df_prior)
    rel_t = np.abs(res.t - eb.t) / np.maximum(1.0, np.abs(eb.t))
    assert rel_t.max() < 1e-2, rel_t.max()
    big = eb.p_value > 1e-250
    dlp = np.abs(np.log(np.maximum(res.P_Value[big], 1e-300)) - np.log(eb.p_value[big]))
    assert dlp.max() < 0.2, ('log p', dlp.max(), res.P_Value[big][np.argmax(dlp)], eb.p_value[big][np.argmax(dlp)])
    # the DEG sets agree up to genes sitting on the thresholds
    a, b = set(res.deg_order.tolist()), set(ref["deg_index"].tolist())
    assert len(a ^ b) <= max(3, 0.02 * len(b)), (len(a), len(b), len(a ^ b))


@pytest.mark.gpu
def test_gpu_lowess_matches_oracle():
    rng = np.random.default_rng(5)
    x = np.sort(np.round(rng.uniform(0, 12, 5000), 2))          # rounding creates ties
    y = 0.3 + 1.5 * np.exp(-0.3 * x) + 0.05 * rng.standard_normal(5000)
    ref = O.lowess(x, y, f=0.5)
    out = DE.lowess_cuda(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), f=0.5)
    assert np.abs(out.cpu().numpy() - ref).max() < 1e-9


@pytest.mark.gpu
def test_gpu_student_t_pvalues():
    from scipy import stats
    from consensusml_b200 import _native
    G = 2000
    rng = np.random.default_rng(6)
    t = np.concatenate([rng.standard_normal(G // 2) * 3, rng.standard_normal(G // 2) * 30])
    coef = torch.from_numpy(t).cuda()
    one = torch.ones(G, dtype=torch.float64, device="cuda")
    outs = [torch.empty(G, dtype=torch.float64, device="cuda") for _ in range(5)]
    P = _native.ptr
    rc = _native.cuda().cml_ebayes(P(coef), P(one), P(one), None, 58.0, G, 1.0, float("inf"), 1e18, 1.0, 0.01,
                                   *[P(o) for o in outs], _native.stream_ptr())
    assert rc == 0
    # infinite prior df -> normal tail
    assert np.allclose(outs[1].cpu().numpy(), 2 * stats.norm.sf(np.abs(t)), rtol=1e-9, atol=1e-300)
    for df0 in (4.0, 300.0):
        rc = _native.cuda().cml_ebayes(P(coef), P(one), P(one), None, 50.0, G, 1.0, df0, 1e18, 1.0, 0.01,
                                       *[P(o) for o in outs], _native.stream_ptr())
        assert rc == 0
        ref = 2 * stats.t.sf(np.abs(t), 50.0 + df0)
        got = outs[1].cpu().numpy()
        ok = ref > 1e-280
        assert np.abs(np.log(got[ok]) - np.log(ref[ok])).max() < 1e-6


def test_sample_sharded_normalisation_and_gene_sharded_voom_two_ranks():
    """stats/sharded.py under gloo with the kernels on the host execution model (tests/dist_stats_check.py)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29552", os.path.join(root, "tests", "dist_stats_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env={**os.environ, "CML_DIST_TIMEOUT_S": "120"})
    assert r.returncode == 0 and "STATS DIST OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
