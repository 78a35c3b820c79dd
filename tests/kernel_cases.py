"""Backend-agnostic checks of the kernels added in round 2.  Each ``case_*`` takes a backend object:

* ``EmuBackend``  - CPU tensors, kernels run on the host execution model (tests/test_emu_kernels.py, part of the CPU suite)
* ``CudaBackend`` - the same assertions on a real GPU (tests/test_y_kernels_gpu.py, ``-m gpu``)

so the kernel source is held to one set of oracles on both.  ``be.dev(t)`` places an input, results are compared on the host."""
import contextlib

import numpy as np
import torch

from consensusml_b200 import _native
from consensusml_b200.io.synthetic import nb_counts


class EmuBackend:
    name = "emu"

    def ctx(self, **kw):
        return _native.emulation(**kw)

    def dev(self, t):
        return t

    def tol(self, emu, cuda):
        return emu


class CudaBackend:
    name = "cuda"

    def ctx(self, **kw):
        return contextlib.nullcontext()

    def dev(self, t):
        return t.cuda()

    def tol(self, emu, cuda):
        return cuda


def _h(t):
    return t.detach().cpu()


def case_column_moments_topk_and_quantile_binning_kernels(be):
    from consensusml_b200.models import trees as TR
    from consensusml_b200.ops import colvar as CV
    g = torch.Generator().manual_seed(12)
    x = torch.randn(300, 70, generator=g) * torch.linspace(0.2, 3.0, 70)[None, :] + 5.0
    x[:, 9] = x[:, 4]                                     # equal variances: tie broken by index
    x[:, 20] = 1.25                                       # constant column
    mask = torch.rand(300, generator=g) < 0.6
    with be.ctx(seed=3):
        cnt, mean, m2 = CV.column_moments(be.dev(x), be.dev(mask))
        idx_hi = _h(CV.variance_topk(be.dev(x), 11, largest=True, row_mask=be.dev(mask)))
        idx_lo = _h(CV.variance_topk(be.dev(x), 7, largest=False))
        tk = _h(CV.topk(be.dev(torch.tensor([3.0, 1.0, 3.0, 2.0, 3.0, -1.0])), 2))
    mean, m2 = _h(mean), _h(m2)
    xs = x[mask].double()
    assert cnt == int(mask.sum())
    assert torch.allclose(mean, xs.mean(0), rtol=1e-12, atol=1e-12)
    assert torch.allclose(m2, ((xs - xs.mean(0)) ** 2).sum(0), rtol=1e-10, atol=1e-10)
    var = xs.var(0, unbiased=True)
    assert idx_hi.tolist() == torch.argsort(var.float(), descending=True, stable=True)[:11].tolist()
    assert idx_lo.tolist() == torch.argsort(x.double().var(0, unbiased=True).float(), descending=False, stable=True)[:7].tolist()
    assert tk.tolist() == [0, 2]
    # quantile binning (n > 256): the kernel path and the host restatement produce the same edges and bins
    xq = torch.randn(700, 9, generator=g)
    xq[:, 2] = torch.round(xq[:, 2] * 2) / 2              # heavy ties -> fewer than max_bin distinct quantiles
    bc, binsc = TR.Binner.fit(xq, max_bin=64)
    with be.ctx(seed=1):
        bd, binsd = TR.Binner.fit(be.dev(xq), max_bin=64)
    assert torch.equal(_h(bd.n_edges), bc.n_edges)
    for f in range(9):
        k = int(bc.n_edges[f])
        assert torch.equal(_h(bd.edges)[f, :k], bc.edges[f, :k]), f
    assert torch.equal(_h(binsd)[:, :700], binsc[:, :700])


def case_kmeans_sweep_kernel_reaches_the_same_minima_as_the_host_path(be):
    from consensusml_b200.models import kmeans as KM
    g = torch.Generator().manual_seed(4)
    centres = torch.tensor([[0.0, 0.0, 0.0], [4.0, 0.0, 1.0], [0.0, 5.0, -1.0], [5.0, 5.0, 3.0]])
    x = torch.cat([c + 0.4 * torch.randn(30, 3, generator=g) for c in centres])
    with be.ctx(seed=2):
        dev = KM.kmeans_sweep(be.dev(x), ks=(4, 3, 2), nstart=10, seed=1000)
        again = KM.kmeans_sweep(be.dev(x), ks=(4, 3, 2), nstart=10, seed=1000)
    for k in (4, 3, 2):
        host = KM.kmeans(x, k, nstart=8, seed=1000)
        r = dev[k]
        r.centers, r.labels, r.withinss = _h(r.centers), _h(r.labels), _h(r.withinss)
        assert r.centers.shape == (k, 3) and r.labels.shape == (120,)
        # k = 4 is the planted structure (one minimum); for k < 4 several local minima lie within a few per cent
        assert abs(r.tot_withinss - host.tot_withinss) <= (1e-3 if k == 4 else 0.03) * host.tot_withinss, (k, r.tot_withinss, host.tot_withinss)
        # internal consistency: centres are the means of their members, withinss adds up, labels are nearest centres
        for j in range(k):
            assert torch.allclose(r.centers[j], x[r.labels == j].mean(0), atol=1e-5)
        assert abs(float(r.withinss.sum()) - r.tot_withinss) < 1e-3
        assert torch.equal(torch.cdist(x, r.centers).argmin(1), r.labels)
        assert torch.equal(r.labels, _h(again[k].labels)) and r.tot_withinss == again[k].tot_withinss      # reproducible
    assert sorted(torch.bincount(dev[4].labels).tolist()) == [30, 30, 30, 30]


def case_batched_metrics_kernel_matches_the_torch_restatement(be):
    from consensusml_b200.ops import metrics as M
    g = torch.Generator().manual_seed(21)
    B, n = 5, 137
    truth = (torch.rand(n, generator=g) < 0.45).float()
    scores = torch.randn(B, n, generator=g) + truth[None, :] * torch.linspace(0, 2, B)[:, None]
    scores[1] = torch.round(scores[1] * 2) / 2            # many tied scores
    scores[4] = 0.3                                       # all tied: AUC 0.5 exactly
    probs = torch.sigmoid(scores)
    with be.ctx(seed=6):
        m = {k: _h(v) for k, v in M.batch_metrics(be.dev(scores), be.dev(truth)).items()}
        mp = {k: _h(v) for k, v in M.batch_metrics(be.dev(probs), be.dev(truth), threshold=0.5, is_prob=True).items()}
        one_auc = M.auc(be.dev(scores[2]), be.dev(truth))
        cm = _h(M.confusion_matrix(be.dev((scores[2] > 0).float()), be.dev(truth)))
        ll = M.log_loss(be.dev(truth), be.dev(probs[3]))
    for b in range(B):
        assert abs(float(m["auc"][b]) - M._auc_torch(scores[b], truth)) < 1e-12, b
        assert abs(float(mp["logloss"][b]) - M._log_loss_torch(truth, probs[b])) < 1e-6
        ref = M._confusion_torch(scores[b] > 0, truth).flatten().tolist()
        assert [int(m[k][b]) for k in ("tn", "fn", "fp", "tp")] == ref
    assert float(m["auc"][4]) == 0.5 and torch.isnan(m["logloss"]).all()
    assert abs(one_auc - M._auc_torch(scores[2], truth)) < 1e-12
    assert torch.equal(cm, M._confusion_torch(scores[2] > 0, truth))
    assert abs(ll - M._log_loss_torch(truth, probs[3])) < 1e-6


def case_glmboost_kernels_follow_the_float64_path(be):
    from consensusml_b200.models import glmboost as GB
    g = torch.Generator().manual_seed(4)
    x = torch.randn(70, 90, generator=g)
    x[:, 11] = 2.0                                         # constant column: never usable
    y = 300 + 40 * x[:, 3] - 25 * x[:, 17] + 5 * torch.randn(70, generator=g)
    ref = GB.glmboost(x, y, mstop=30, nu=0.1)
    with be.ctx(seed=3):
        fit = GB.glmboost(be.dev(x), be.dev(y), mstop=30, nu=0.1)
    assert fit.path == ref.path
    assert torch.allclose(_h(fit.coef), ref.coef, rtol=1e-5, atol=1e-6) and abs(fit.intercept - ref.intercept) < 1e-3
    assert np.allclose(fit.risk, ref.risk, rtol=1e-5)


def case_mlp_epoch_kernels_match_a_torch_restatement_with_the_same_bf16_roundings(be):
    from consensusml_b200.models import mlp as MLP
    g = torch.Generator().manual_seed(5)
    n, p = 61, 37
    x = torch.randn(n, p, generator=g)
    y = ((x[:, 0] - x[:, 1] + 0.5 * x[:, 2]) > 0).float()
    grid = [dict(hidden=5, lr=3e-3, weight_decay=0.0, seed=0), dict(hidden=9, lr=1e-2, weight_decay=1e-3, seed=1),
            dict(hidden=3, lr=1e-3, weight_decay=1e-2, seed=2)]
    # one epoch must agree to rounding; over more epochs Adam's m / sqrt(v) amplifies last-bit differences of tiny
    # gradients (and a bf16 rounding tie can flip), so the 12-epoch comparison is looser
    # (on a GPU the two products are tcgen05 GEMMs: fp32 accumulation in a different order than torch's)
    for epochs, tol in ((1, be.tol(2e-6, 2e-5)), (12, be.tol(1e-3, 3e-3))):
        _check_mlp(be, MLP, x, y, grid, n, epochs, tol)


def _check_mlp(be, MLP, x, y, grid, n, epochs, tol):
    with be.ctx(seed=4):
        res = MLP.train_grid(be.dev(x), be.dev(y), grid, epochs=epochs)
    # restatement: the fp32 torch loop of train_grid with bf16 rounding of the three GEMM operands (X, W1, dH)
    bf = lambda t: t.to(torch.bfloat16).float()
    xf = x.float()
    mean, sd = xf.mean(0), xf.std(0, unbiased=True)
    xs = (xf - mean) / torch.where(sd > 0, sd, torch.ones_like(sd))
    ref = MLP.train_grid(x, y, grid, epochs=0)                              # initial weights only
    w1, b1, w2, b2 = ref.w1.clone(), ref.b1.clone(), ref.w2.clone(), ref.b2.clone()
    hs = [c["hidden"] for c in grid]
    seg = torch.repeat_interleave(torch.arange(3), torch.tensor(hs))
    lr_u = torch.tensor([c["lr"] for c in grid])[seg]
    wd_u = torch.tensor([c["weight_decay"] for c in grid])[seg]
    lr_g = torch.tensor([c["lr"] for c in grid])
    params = [w1, b1, w2, b2]
    m = [torch.zeros_like(t) for t in params]
    v = [torch.zeros_like(t) for t in params]
    lrs, wds = [lr_u[:, None], lr_u, lr_u, lr_g], [wd_u[:, None], None, wd_u, None]
    xb = bf(xs)
    for t in range(1, epochs + 1):
        pre = xb @ bf(w1).t() + b1
        h = torch.relu(pre)
        logit = torch.zeros(n, 3).index_add_(1, seg, h * w2[None, :]) + b2[None, :]
        prob = torch.sigmoid(logit)
        dlogit = (prob - y[:, None]) / n
        dh = dlogit[:, seg] * w2[None, :] * (pre > 0)
        grads = [bf(dh).t() @ xb, dh.sum(0), (dlogit[:, seg] * h).sum(0), dlogit.sum(0)]
        for i, (pt, gr) in enumerate(zip(params, grads)):
            if wds[i] is not None:
                gr = gr + wds[i] * pt
            m[i].mul_(0.9).add_(gr, alpha=0.1)
            v[i].mul_(0.999).addcmul_(gr, gr, value=0.001)
            pt.sub_(lrs[i] * (m[i] / (1 - 0.9 ** t)) / ((v[i] / (1 - 0.999 ** t)).sqrt() + 1e-8))
    for a, b in ((res.w1, w1), (res.b1, b1), (res.w2, w2), (res.b2, b2)):
        assert float((_h(a) - b).abs().max()) <= tol, (epochs, float((_h(a) - b).abs().max()))
    assert torch.isfinite(_h(res.train_loss)).all() and (_h(res.train_loss) < 0.8).all()


def _gbm_problem(n=900, p=21, seed=3):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, p, generator=g)
    beta = torch.zeros(p)
    beta[:6] = torch.tensor([1.2, -1.0, 0.8, -0.8, 0.6, 0.5])
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ beta + 0.8 * x[:, 6] * x[:, 7])).float()
    return x, y


def case_large_gbm_kernel_level_loop_builds_the_same_trees_as_the_torch_path(be):
    """Histogram build (slot table, reduce-scatter landing with world = 1), sibling derivation, split scoring, totals and
    routing kernels against the torch restatement of the same level-wise algorithm."""
    from consensusml_b200.models.gbm_large import GBMLarge, quantile_bin
    from consensusml_b200.ops.metrics import _auc_torch
    x, y = _gbm_problem()
    bins, cuts = quantile_bin(x, 32)
    ref = GBMLarge(n_estimators=3, max_depth=3, nbins=32, backend="unfused", sibling_subtraction=True).fit(None, y, bins=bins, cuts=cuts)
    for sib in (True, False):
        with be.ctx(seed=5, sms=2):
            got = GBMLarge(n_estimators=3, max_depth=3, nbins=32, sibling_subtraction=sib).fit(None, be.dev(y), bins=be.dev(bins), cuts=be.dev(cuts))
        for a, b in zip(ref.trees, got.trees):
            assert torch.equal(a.feat, _h(b.feat)), (sib, a.feat, b.feat)
            split = a.feat >= 0
            assert torch.equal(a.thr[split], _h(b.thr)[split])
            assert (a.value - _h(b.value)).abs().max() < 2e-4
        assert torch.allclose(ref.importance_gain_, _h(got.importance_gain_), rtol=2e-3, atol=1e-3)
    with be.ctx(seed=1):
        pr_kernel = _h(got.predict_proba(be.dev(x)))                     # gbm_predict_kernel
    got_host = GBMLarge(max_depth=3, nbins=32)
    got_host.trees = [type(t)(_h(t.feat), _h(t.thr), _h(t.value)) for t in got.trees]
    got_host.cuts = _h(got.cuts)
    assert torch.allclose(pr_kernel, got_host.predict_proba(x), atol=1e-6)      # ... equals the torch walk over the same trees
    assert _auc_torch(pr_kernel, y) > 0.8


def case_size_factor_and_plda_score_kernels(be):
    from consensusml_b200.models import discriminant as D
    c, y = nb_counts(120, 31, seed=8, n_de=30)
    c = torch.from_numpy(c) + 1
    c[5, 3] = 0                                           # a gene with a zero is skipped by the median of ratios
    yt = torch.from_numpy(y).long()
    sf_ref = D.deseq_size_factors(c)
    m_ref = D.PoissonLDA(rho=2.0).fit(c, yt)
    m2_ref = D.PoissonLDA(power=True).fit(c, yt)
    with be.ctx(seed=3):
        sf = _h(D.deseq_size_factors(be.dev(c)))
        m_dev = D.PoissonLDA(rho=2.0).fit(be.dev(c), be.dev(yt))
        sc = _h(m_dev.scores(be.dev(c)))
        m2_dev = D.PoissonLDA(power=True).fit(be.dev(c), be.dev(yt))
        sc2 = _h(m2_dev.scores(be.dev(c[:, :7].contiguous())))
    assert torch.allclose(sf, sf_ref, rtol=1e-12, atol=1e-12)
    assert torch.allclose(sc, m_ref.scores(c), rtol=1e-9, atol=1e-6)
    assert torch.allclose(sc2, m2_ref.scores(c[:, :7].contiguous()), rtol=1e-9, atol=1e-6)


def case_large_svm_subproblems_on_the_warm_started_smo_kernel(be):
    """Working-set decomposition (models/svm_large.py) with the sub-problems solved by ``cml_svm_smo_warm``."""
    from consensusml_b200.models.svm_large import LargeSVC
    g = torch.Generator().manual_seed(3)
    n, p = 160, 6
    y = (torch.rand(n, generator=g) < 0.5).float()
    x = torch.randn(n, p, generator=g) + (y[:, None] - 0.5) * torch.tensor([1.6, -1.2, 0.8, 0.0, 0.0, 0.0])
    xs = (x - x.mean(0)) / x.std(0, unbiased=True)
    ref = LargeSVC(kernel="radial", q=32, tol=1e-3).fit(xs, y)                       # torch / NumPy path
    with be.ctx(seed=2):
        got = LargeSVC(kernel="radial", q=32, tol=1e-3).fit(be.dev(xs), be.dev(y))
    # (on a GPU the kernel columns come from the bf16 tensor-core Gram kernel: a slightly different kernel matrix)
    assert abs(got.objective_ - ref.objective_) < be.tol(2e-3, 2e-2) * max(1.0, abs(ref.objective_))
    assert got.gap_ < 1e-3 and abs(got.rho_ - ref.rho_) < be.tol(2e-2, 5e-2)


def case_lasso_cta_per_fit_kernel_is_bit_identical_to_the_warp_kernel(be):
    """Latency mode (one CTA per fit: shared-memory column cache, eight warps on the p-wide passes) against the validated
    one-warp kernel: same arithmetic, so every output must be equal bit for bit, in all three solver variants."""
    import os
    from consensusml_b200.ops import lasso as K
    g = torch.Generator().manual_seed(3)
    n, p, B = 75, 200, 3
    x = torch.randn(n, p, generator=g)
    beta = torch.zeros(p)
    beta[:6] = torch.tensor([1.0, -1.0, 0.8, -0.8, 0.5, 0.5])
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ beta)).float()
    w = torch.ones(B, n)
    for b in range(1, B):
        w[b, torch.randperm(n, generator=g)[:7]] = 0
    mask = torch.ones(B, p, dtype=torch.bool)
    mask[1, :4] = False
    lam = np.geomspace(0.3, 0.02, 14)
    old = os.environ.get("CML_LASSO_TEAM")
    try:
        # the last variant shrinks the column cache to 5 slots: most screened features then stay in global memory and the
        # solver mixes cached and uncached coordinates
        for kw in (dict(), dict(alpha=0.5, intercept=True), dict(modified_newton=True), dict(_cap="5")):
            kw = dict(kw)
            cap = kw.pop("_cap", None)
            if cap is None:
                os.environ.pop("CML_LASSO_TEAM_CAP", None)
            else:
                os.environ["CML_LASSO_TEAM_CAP"] = cap
            res = {}
            for mode in ("0", "1"):
                os.environ["CML_LASSO_TEAM"] = mode
                with be.ctx(seed=int(mode) * 3 + 1):
                    freq = be.dev(torch.zeros(len(lam), p))
                    o = K.lognet_path(be.dev(x), be.dev(y), be.dev(w), lam, feature_mask=be.dev(mask), freq_out=freq, **kw)
                    res[mode] = (o, _h(freq))
            a, fa = res["0"]
            b_, fb = res["1"]
            assert int(_h(a.jerr).abs().max()) == 0
            for name in ("npasses", "kin", "eta", "a0", "dev_ratio", "nulldev", "lmu"):
                assert torch.equal(_h(getattr(a, name)), _h(getattr(b_, name))), (kw, name)
            assert torch.equal(fa, fb)
            for i in range(B):
                assert torch.equal(_h(a.dense(i)), _h(b_.dense(i))), (kw, i)
    finally:
        os.environ.pop("CML_LASSO_TEAM_CAP", None)
        if old is None:
            os.environ.pop("CML_LASSO_TEAM", None)
        else:
            os.environ["CML_LASSO_TEAM"] = old


def case_lasso_cta_per_fit_kernel_error_paths_match_the_warp_kernel(be):
    """Active-list overflow, a replicate without weight, a one-class replicate and the iteration cap: the paths on which the
    solver warp leaves its loops early while the helper warps are parked at a barrier."""
    import os
    from consensusml_b200.ops import lasso as K
    g = torch.Generator().manual_seed(5)
    n, p, B = 40, 90, 4
    x = torch.randn(n, p, generator=g)
    y = (x[:, 0] - x[:, 1] + 0.5 * torch.randn(n, generator=g) > 0).float()
    w = torch.ones(B, n)
    w[1] = 0
    yy = torch.stack([y, y, torch.ones_like(y), y])
    lam = np.geomspace(0.3, 0.02, 14)
    old = os.environ.get("CML_LASSO_TEAM")
    res = {}
    try:
        for mode in ("0", "1"):
            os.environ["CML_LASSO_TEAM"] = mode
            with be.ctx(seed=2):
                res[mode] = (K.lognet_path(be.dev(x), be.dev(y), be.dev(w), lam, nx=6),
                             K.lognet_path(be.dev(x), be.dev(yy), be.dev(w), lam, maxit=40))
    finally:
        if old is None:
            os.environ.pop("CML_LASSO_TEAM", None)
        else:
            os.environ["CML_LASSO_TEAM"] = old
    for i in (0, 1):
        a, b_ = res["0"][i], res["1"][i]
        for name in ("jerr", "lmu", "npasses", "kin"):
            assert torch.equal(_h(getattr(a, name)), _h(getattr(b_, name))), (i, name)
    j0, j1 = _h(res["1"][0].jerr).tolist(), _h(res["1"][1].jerr).tolist()
    assert j0[1] == 10000 and all(v < -10000 for v in (j0[0], j0[2], j0[3]))          # no weight; active-list overflow
    assert j1[1] == 10000 and j1[2] == 9001 and j1[0] < 0 and j1[3] < 0                # one class; iteration cap


def case_forest_proximity_as_one_hot_gemm(be):
    """randomForest-style proximity through the one-hot membership kernel + GEMM against the definition."""
    from consensusml_b200.models import trees as TR
    g = torch.Generator().manual_seed(1)
    T, n, max_nodes = 37, 70, 9
    leaf = torch.randint(0, max_nodes, (T, n), generator=g, dtype=torch.int32)
    ob = (torch.rand(T, n, generator=g) < 0.4).to(torch.uint8)
    both = (ob[:, :, None] & ob[:, None, :]).bool()
    want_p = ((leaf[:, :, None] == leaf[:, None, :]) & both).float().sum(0)
    with be.ctx(seed=4):
        p1, q1 = TR.proximity_gemm(be.dev(leaf), be.dev(ob), max_nodes, max_cols=100)        # several tree groups
        p2, q2 = TR.proximity_gemm(be.dev(leaf), None, max_nodes)
    assert torch.equal(_h(p1), want_p) and torch.equal(_h(q1), both.float().sum(0))
    assert torch.equal(_h(p2), (leaf[:, :, None] == leaf[:, None, :]).float().sum(0)) and float(_h(q2).min()) == T


def case_deseq_dispersion_and_rlog_kernels_follow_the_float64_path(be):
    """csrc/deseq.cu against the vectorised float64 restatement of the same DESeq2 steps (models/deseq.py)."""
    from consensusml_b200.models import deseq as DQ
    rng = np.random.default_rng(31)
    G, S = 180, 45
    mu = np.exp(rng.normal(4, 1.5, G))
    disp = 0.05 + 2.0 / mu
    sf0 = np.exp(rng.normal(0, 0.3, S))
    c = torch.from_numpy(rng.negative_binomial(1 / disp[:, None], 1 / (1 + disp[:, None] * mu[:, None] * sf0[None, :])).astype(np.float64))
    c[3] = 0
    groups = torch.from_numpy(np.arange(S) % 3)
    ref = DQ.estimate_dispersions(c, groups=groups)                 # CPU tensors outside emulation: the torch path
    rr = DQ.rlog(c, ref)
    fr = DQ.rlog(c[:, :6], frozen=rr)
    with be.ctx(seed=5):
        got = DQ.estimate_dispersions(be.dev(c), groups=be.dev(groups))
        gr = DQ.rlog(be.dev(c), got)
        gf = DQ.rlog(be.dev(c[:, :6]), frozen=gr)
    ok = torch.isfinite(ref.disp_gene)
    assert torch.equal(torch.isfinite(_h(got.disp_gene)), ok)
    di = (_h(got.iters) - ref.iters).abs()
    assert int(di.max()) <= 2 and float((di == 0).float().mean()) > 0.9
    rel = (_h(got.disp_gene)[ok] - ref.disp_gene[ok]).abs() / ref.disp_gene[ok]
    assert float(rel.max()) < 5e-3 and float(rel.median()) < 1e-6
    assert got.fit_type == ref.fit_type and np.allclose(got.coefs, ref.coefs, rtol=1e-3)
    assert (_h(gr.values) - rr.values).abs().max() < 2e-3 and (_h(gf.values) - fr.values).abs().max() < 2e-3
    assert abs(gr.beta_prior_var - rr.beta_prior_var) < 1e-3 * rr.beta_prior_var


def case_voom_wls_wide_design_on_the_lane_distributed_inverse(be):
    from consensusml_b200.stats import de as DE
    c, g = nb_counts(300, 36, seed=26, n_de=40)
    rng = np.random.default_rng(6)
    k = 6
    X = np.column_stack([np.ones(36), np.asarray(g, dtype=np.float64)] + [rng.normal(size=36) for _ in range(k - 2)])
    cvec = np.zeros(k)
    cvec[1] = 1.0
    ref = DE.voom_de(torch.from_numpy(c), g, design=X, contrast=cvec)
    with be.ctx(seed=2):
        res = DE.voom_de(be.dev(torch.from_numpy(c)), g, design=X, contrast=cvec)
    assert np.array_equal(np.asarray(_h(torch.as_tensor(res.keep))), np.asarray(ref.keep))
    assert np.abs(np.asarray(res.logFC) - np.asarray(ref.logFC)).max() < 2e-3
    t_a, t_b = np.asarray(res.t), np.asarray(ref.t)
    assert (np.abs(t_a - t_b) / np.maximum(1.0, np.abs(t_b))).max() < 1e-2


def case_batched_sharded_lasso_iteration_on_the_tensor_core_kernels(be):
    """BatchedLassoLarge on one rank: tcgen05 eta product, residual split, fused GEMM -> reduce-scatter, proximal step + all-gather."""
    from consensusml_b200.models import lasso_large as LL
    g = torch.Generator().manual_seed(4)
    n, p, B = 140, 200, 10
    x = torch.randn(n, p, generator=g).bfloat16().float()
    bt = torch.zeros(p)
    bt[:6] = torch.randn(6, generator=g) * 1.5
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ bt)).float()
    lmax = LL.lambda_max(x, y, n)
    lams = np.geomspace(lmax * 0.9, lmax * 0.1, B)
    W = torch.poisson(torch.ones(B, n), generator=g)
    ref = LL.fit_batched(x, y, lams, weights=W, backend="unfused", max_iter=30, tol=0.0)
    with be.ctx(seed=3):
        got = LL.BatchedLassoLarge(be.dev(x), be.dev(y), lams, weights=be.dev(W), step=ref.step).run(30, 0.0, 10)
    assert (_h(got.beta) - ref.beta).abs().max() < be.tol(2e-4, 5e-4)
    assert (got.objective - ref.objective).abs().max() < 1e-4


def case_lasso_path_results_do_not_depend_on_the_launch(be):
    """The same batch three times (on the host model: under three thread orders / shared-memory garbage patterns): pass counts,
    active sets and coefficients must be EQUAL - the round-1 kernel differed "by a sweep or two" on a first launch because its
    convergence test read a shared-memory slot nobody had written."""
    from consensusml_b200.ops import lasso as K
    g = torch.Generator().manual_seed(8)
    n, p, B = 75, 300, 6
    x = torch.randn(n, p, generator=g)
    bt = torch.zeros(p)
    bt[:6] = torch.tensor([1.0, -1.0, 0.8, -0.8, 0.5, 0.5])
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ bt)).float()
    w = torch.ones(B, n)
    for b in range(1, B):
        w[b, torch.randperm(n, generator=g)[:7]] = 0
    lam = 10.0 ** np.linspace(-0.3, -2.0, 40)
    runs = []
    for seed in (1, 2, 3):
        with be.ctx(seed=seed):
            pb = K.lognet_path(be.dev(x), be.dev(y), be.dev(w), lam)
            runs.append((_h(pb.npasses).clone(), _h(pb.kin).clone(), _h(pb.dense(0)).clone(), _h(pb.dense(B - 1)).clone()))
    for r in runs[1:]:
        for a, b in zip(runs[0], r):
            assert torch.equal(a, b)
    assert int(runs[0][0].min()) > 40


def case_shallow_trees_grown_without_host_syncs_equal_the_synchronised_growth(be):
    """grow() for max_depth <= 8 launches every level with an upper bound on the frontier and reads the true size on the device;
    the trees (boosting rounds in both modes, a depth-3 forest) must be the ones the level-by-level synchronised loop grows."""
    from consensusml_b200.models import trees as TR
    g = torch.Generator().manual_seed(6)
    n, p = 120, 40
    x = torch.randn(n, p, generator=g)
    y = ((x[:, 0] - 0.7 * x[:, 1] + 0.4 * torch.randn(n, generator=g)) > 0).float()

    def fit_all():
        a = TR.GradientBoosting(n_estimators=6, learning_rate=0.3, max_depth=3, mode="xgboost", seed=8).fit(be.dev(x), be.dev(y))
        b = TR.GradientBoosting(n_estimators=6, learning_rate=0.1, max_depth=4, max_features="log2", min_samples_leaf=4, subsample=0.5, mode="sklearn",
                                seed=8).fit(be.dev(x), be.dev(y))
        c = TR.RandomForest(n_trees=20, max_depth=3, seed=8).fit(be.dev(x), be.dev(y))
        d = TR.RandomForest(n_trees=12, seed=8).fit(be.dev(x), be.dev(y))                       # unlimited depth: bounded levels first, then the compacted loop
        e = TR.GradientBoosting(n_estimators=3, learning_rate=0.3, max_depth=12, mode="xgboost", seed=8).fit(be.dev(x), be.dev(y))
        return a, b, c, d, e
    with be.ctx(seed=4):
        _native.reset_calls()
        fast = fit_all()
        assert _native.calls.get("cml_forest_level_dev", 0) > 0 and _native.calls.get("cml_forest_level", 0) > 0
        old = TR.SYNC_FREE_MAX_DEPTH
        TR.SYNC_FREE_MAX_DEPTH = -1
        try:
            slow = fit_all()
        finally:
            TR.SYNC_FREE_MAX_DEPTH = old
    for f, s in zip(fast[:2], slow[:2]):
        assert len(f.trees) == len(s.trees)
        for tf, ts in zip(f.trees, s.trees):
            if be.name == "emu":       # (on a GPU the children of one level get their node ids in atomic order: compare what the trees compute)
                assert torch.equal(_h(tf.feat), _h(ts.feat)) and torch.equal(_h(tf.thr), _h(ts.thr)) and torch.equal(_h(tf.left), _h(ts.left))
                assert torch.allclose(_h(tf.value), _h(ts.value), atol=1e-6)
            assert int(_h(tf.node_count).sum()) == int(_h(ts.node_count).sum())
        assert torch.allclose(_h(f.predict_proba(be.dev(x))), _h(s.predict_proba(be.dev(x))), atol=1e-6)
    for k in (2, 3):
        assert abs(fast[k].oob_error_ - slow[k].oob_error_) < 1e-9
        assert torch.allclose(_h(fast[k].importance_), _h(slow[k].importance_), atol=1e-6)
    assert torch.allclose(_h(fast[4].predict_proba(be.dev(x))), _h(slow[4].predict_proba(be.dev(x))), atol=1e-6)
