"""The CUDA kernels, run on the CPU through the host execution model (csrc/emu/cuda_emu.h).

``libcml_emu.so`` is the SAME kernel source as ``libcml_b200.so`` (csrc/*.cu minus the tcgen05 GEMMs) compiled by g++:
every CUDA thread is a fiber, warp shuffles / ballots / barriers are rendezvous points, shared memory starts as
seed-dependent garbage.  These tests call the public op wrappers with CPU tensors inside ``_native.emulation()`` and
hold the kernels to the float64 oracles - the same comparisons the ``-m gpu`` tests make on hardware, at smaller sizes -
so a logic error, an uninitialised read or a missing ``__syncwarp`` is caught on a machine without a GPU.
(The reference has no tests and no native code)"""
import ctypes

import numpy as np
import pytest
import torch

from consensusml_b200 import _native
from consensusml_b200.io.synthetic import nb_counts
from consensusml_b200.oracle import limma as O
from consensusml_b200.ops import lasso as K


def _lasso_problem(seed=0, n=75, p=200, B=4):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, p, generator=g)
    beta = torch.zeros(p)
    beta[:6] = torch.tensor([1.0, -1.0, 0.8, -0.8, 0.5, 0.5])
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ beta)).float()
    w = torch.ones(B, n)
    for b in range(1, B):
        w[b, torch.randperm(n, generator=g)[:7]] = 0
    return x, y, w, np.geomspace(0.3, 0.02, 14)


def test_lasso_path_kernel_matches_host_oracle_statement_by_statement():
    x, y, w, lam = _lasso_problem()
    for kw in (dict(), dict(alpha=0.5, intercept=True), dict(modified_newton=True)):
        host = K.lognet_path(x.double(), y.double(), w.double(), lam, **kw)
        with _native.emulation():
            dev = K.lognet_path(x, y, w, lam, **kw)
        assert (dev.jerr == 0).all()
        # the kernel follows the host solver's update order, so even the pass counts agree (fp32 vs fp64 can move a
        # threshold decision by a sweep at most)
        assert (dev.npasses - host.npasses).abs().max() <= 2, (kw, dev.npasses, host.npasses)
        for b in range(w.shape[0]):
            assert (host.dense(b) - dev.dense(b).double()).abs().max() < 1e-4
        assert (host.eta - dev.eta.double()).abs().max() < 1e-3
        assert (host.dev_ratio - dev.dev_ratio.double()).abs().max() < 1e-4


def test_lasso_path_kernel_does_not_depend_on_stale_shared_memory_or_thread_order(monkeypatch):
    """Regression test for the launch-to-launch variation of round 1: the convergence test read the IRLS-start
    coefficient of a feature that had only just entered the active list from a shared-memory slot nobody had written.
    With NaN garbage there the kernel declared convergence one iteration early."""
    x, y, w, lam = _lasso_problem(seed=3)
    runs = []
    for seed, poison in ((0, None), (7, None), (0, "ff"), (0, "00")):
        if poison:
            monkeypatch.setenv("CML_EMU_POISON", poison)
        else:
            monkeypatch.delenv("CML_EMU_POISON", raising=False)
        with _native.emulation(seed=seed):
            runs.append(K.lognet_path(x, y, w, lam))
    ref = runs[0]
    for o in runs[1:]:
        assert torch.equal(o.npasses, ref.npasses)
        assert torch.equal(o.kin, ref.kin) and torch.equal(o.eta, ref.eta) and torch.equal(o.a0, ref.a0)
        for b in range(w.shape[0]):
            assert torch.equal(o.dense(b), ref.dense(b))


def test_lasso_static_scratch_entry_is_bit_identical_to_the_regular_entry():
    import test_zz_no_alloc as T
    with _native.emulation():
        lib = _native.cuda()
        T._declare(lib)
        x, y, w, lam = T._problem(5, 5, 40, 90, 6)
        ref = T._run(lib, False, x, y, w, lam, nx=32, device="cpu")
        got = T._run(lib, True, x, y, w, lam, nx=32, device="cpu")
        again = T._run(lib, True, x, y, w, lam, nx=32, device="cpu")
    for k in ("kin", "npasses", "a0", "dev", "eta", "freq", "jerr", "lmu"):
        assert torch.equal(ref[k], got[k]) and torch.equal(ref[k], again[k]), k
    for b in range(5):
        assert torch.equal(T._dense(ref, b, 90), T._dense(got, b, 90))


def test_normalisation_kernels_match_oracle():
    from consensusml_b200.stats import normalize as N
    c, _ = nb_counts(700, 14, seed=3, n_de=60)
    ct = torch.from_numpy(c)
    with _native.emulation(seed=2):
        lib = N.lib_sizes(ct)
        keep = N.cpm_filter(ct, 1.0, 5, lib=lib)
        sub = ct[keep].contiguous()
        f = N.tmm_factors(sub, lib)
        lc = N.log_cpm(sub, lib, f)
        lct = N.log_cpm(sub, lib, f, sample_major=True)
    assert np.array_equal(lib.numpy(), c.sum(0).astype(float))
    keep_ref = O.cpm_filter(c, 1.0, 5)
    assert np.array_equal(keep.numpy(), keep_ref)
    f_ref = O.tmm_factors(c[keep_ref].astype(float), c.sum(0).astype(float))
    assert np.abs(f.numpy() - f_ref).max() < 2e-4
    lc_ref = O.log_cpm(c[keep_ref].astype(float), c.sum(0).astype(float), f.numpy())
    assert np.abs(lc.numpy() - lc_ref).max() < 1e-4
    assert torch.equal(lct.t().contiguous(), lc)


def test_tmm_kernel_with_ties_and_zeros():
    from consensusml_b200.stats import normalize as N
    rng = np.random.default_rng(9)
    c = rng.poisson(3, (900, 6)).astype(np.int32)
    c[:, 3] *= 2
    with _native.emulation():
        f = N.tmm_factors(torch.from_numpy(c)).numpy()
    assert np.abs(f - O.tmm_factors(c.astype(float))).max() < 5e-4


def test_voom_de_kernels_match_oracle():
    from consensusml_b200.stats import de as DE
    c, g = nb_counts(900, 24, seed=11, n_de=90)
    ref = O.voom_de(c, g)
    with _native.emulation(seed=1):
        res = DE.voom_de(torch.from_numpy(c), g)
    eb = ref["ebayes"]
    assert np.array_equal(res.keep.numpy(), ref["keep"])
    assert np.abs(res.logFC - ref["coef"]).max() < 2e-3
    assert np.abs(res.AveExpr - ref["fit"].amean).max() < 1e-3
    assert abs(res.s2_prior - eb.s2_prior) / eb.s2_prior < 5e-3
    assert abs(res.df_prior - eb.df_prior) / eb.df_prior < 2e-2
    assert (np.abs(res.t - eb.t) / np.maximum(1.0, np.abs(eb.t))).max() < 1e-2
    a, b = set(res.deg_order.tolist()), set(ref["deg_index"].tolist())
    assert len(a ^ b) <= max(3, 0.03 * len(b))


def test_lowess_kernel_matches_oracle():
    from consensusml_b200.stats import de as DE
    rng = np.random.default_rng(5)
    x = np.sort(np.round(rng.uniform(0, 12, 400), 2))
    y = 0.3 + 1.5 * np.exp(-0.3 * x) + 0.05 * rng.standard_normal(400)
    with _native.emulation():
        out = DE.lowess_cuda(torch.from_numpy(x), torch.from_numpy(y), f=0.5)
    assert np.abs(out.numpy() - O.lowess(x, y, f=0.5)).max() < 1e-9


def test_student_t_tail_kernel():
    from scipy import stats
    from consensusml_b200.stats import de as DE   # noqa: F401  (registers cml_ebayes)
    G = 300
    rng = np.random.default_rng(6)
    t = np.concatenate([rng.standard_normal(G // 2) * 3, rng.standard_normal(G // 2) * 30])
    coef, one = torch.from_numpy(t), torch.ones(G, dtype=torch.float64)
    outs = [torch.empty(G, dtype=torch.float64) for _ in range(5)]
    P = _native.ptr
    with _native.emulation():
        for df0 in (4.0, 300.0):
            rc = _native.cuda().cml_ebayes(P(coef), P(one), P(one), None, 50.0, G, 1.0, df0, 1e18, 1.0, 0.01,
                                           *[P(o) for o in outs], _native.stream_ptr())
            assert rc == 0
            ref = 2 * stats.t.sf(np.abs(t), 50.0 + df0)
            ok = ref > 1e-280
            assert np.abs(np.log(outs[1].numpy()[ok]) - np.log(ref[ok])).max() < 1e-6


def _tree_data(n=60, p=40, seed=1):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, p, generator=g)
    x[:, 5] = torch.round(x[:, 5])                    # ties
    y = ((x[:, 0] + 0.7 * x[:, 3] - 0.5 * x[:, 5] + 0.3 * torch.randn(n, generator=g)) > 0).float()
    return x, y


def test_binning_and_tree_kernels_match_numpy_restatement():
    from consensusml_b200.models import trees as TR
    from consensusml_b200.ops.replicate import replicate_weights
    x, y = _tree_data()
    n = x.shape[0]
    bc, binsc = TR.Binner.fit(x)
    with _native.emulation(seed=3):
        bg, binsg = TR.Binner.fit(x)
        tr_new = bg.transform(x[:17] + 0.01)
    assert torch.equal(binsg[:, :n], binsc[:, :n]) and torch.equal(bg.n_edges, bc.n_edges)
    assert torch.equal(tr_new[:, :17], bc.transform(x[:17] + 0.01)[:, :17])
    w = replicate_weights(5, n, seed=5)[:, :n].contiguous()
    kw = dict(criterion=TR.GINI, mtry=6, max_bin=bc.max_bin, seed=11, in_tree=w > 0)
    cpu = TR.grow(binsc, n, w * y[None, :], w, **kw)
    with _native.emulation(seed=4):
        dev = TR.grow(binsc, n, w * y[None, :], w, **kw)
        ld, vd = TR.predict_leaves(dev, binsc, n)
    assert torch.equal(dev.node_count, cpu.node_count)

    def splits(tb, t):
        nc = int(tb.node_count[t])
        return sorted((int(a), int(b), round(float(c), 3)) for a, b, c in zip(tb.feat[t, :nc], tb.thr[t, :nc], tb.cover[t, :nc]) if a >= 0)
    for t in range(5):
        assert splits(dev, t) == splits(cpu, t), t
    lc, vc = TR.predict_leaves(cpu, binsc, n)
    assert torch.allclose(vd, vc, atol=1e-6)
    for t in range(5):
        a, b = lc[t], ld[t]
        assert torch.equal(a[:, None] == a[None, :], b[:, None] == b[None, :])


def test_forest_and_boosting_through_the_kernels():
    from consensusml_b200.models import trees as TR
    from consensusml_b200.ops import metrics as M
    x, y = _tree_data(n=80, p=30, seed=4)
    with _native.emulation():
        rf = TR.RandomForest(n_trees=40, proximity=True, seed=3).fit(x, y)
        gb = TR.GradientBoosting(n_estimators=8, learning_rate=0.5, max_depth=2, mode="xgboost").fit(x, y)
        pr = gb.predict_proba(x)
    # bootstrap rows come from the replicate kernel under emulation and from NumPy otherwise (same distribution, different
    # stream), so the forests are compared statistically; the boosted trees below use no randomness and must agree
    rf_cpu = TR.RandomForest(n_trees=40, proximity=True, seed=3).fit(x, y)
    assert abs(rf.oob_error_ - rf_cpu.oob_error_) < 0.12 and rf.oob_error_ < 0.4
    assert set(torch.argsort(-rf.importance_)[:2].tolist()) & {0, 3, 5}
    assert abs(float(rf.importance_.sum()) - float(rf_cpu.importance_.sum())) / float(rf_cpu.importance_.sum()) < 0.2
    pm = rf.proximity_
    assert torch.allclose(pm, pm.t()) and float(pm.min()) >= 0 and float(pm.max()) <= 1 + 1e-6
    # the proximity kernel itself, on fixed inputs, against the definition
    g = torch.Generator().manual_seed(1)
    leaf = torch.randint(0, 4, (12, 21), generator=g, dtype=torch.int32)
    ob = (torch.rand(12, 21, generator=g) < 0.5).to(torch.uint8)
    prox, pairs = torch.zeros(21, 21), torch.zeros(21, 21)
    with _native.emulation(seed=9):
        rc = _native.cuda().cml_forest_proximity(_native.ptr(leaf), _native.ptr(ob), 12, 21, _native.ptr(prox), _native.ptr(pairs), None)
    assert rc == 0
    both = (ob[:, :, None] & ob[:, None, :]).bool()
    assert torch.equal(prox, ((leaf[:, :, None] == leaf[:, None, :]) & both).float().sum(0))
    assert torch.equal(pairs, both.float().sum(0))
    gb_cpu = TR.GradientBoosting(n_estimators=8, learning_rate=0.5, max_depth=2, mode="xgboost").fit(x, y)
    assert np.allclose(gb.train_error_, gb_cpu.train_error_)
    assert M.auc(pr, y) > 0.9


def test_smo_kernel_matches_numpy_smo():
    from consensusml_b200.models import svm as S
    g = torch.Generator().manual_seed(2)
    n = 50
    x = torch.randn(n, 8, generator=g)
    y = torch.where(x[:, 0] + 0.5 * x[:, 1] + 0.3 * torch.randn(n, generator=g) > 0, 1.0, -1.0)
    Km = torch.exp(-0.1 * torch.cdist(x, x) ** 2)
    cb = torch.ones(2, n)
    cb[1, :5] = 0                                      # second problem leaves five samples out
    a_ref, rho_ref, _ = S.solve_dual(Km.double(), y.double(), cb.double())
    with _native.emulation(seed=5):
        a, rho, it = S.solve_dual(Km, y, cb)
    def obj(al):
        ay = al.double() * y.double()
        return 0.5 * (ay @ Km.double() @ ay) - al.double().sum()
    for b in range(2):
        assert abs(float(obj(a[b]) - obj(a_ref[b]))) < 2e-3 * max(1.0, abs(float(obj(a_ref[b]))))
        assert abs(float(rho[b]) - float(rho_ref[b])) < 2e-2
        assert float(a[b][cb[b] == 0].abs().max() if (cb[b] == 0).any() else 0.0) == 0.0


def test_rank_standardise_kde_and_assign_kernels():
    from consensusml_b200.stats import correlation as C
    import scipy.stats as sps
    g = torch.Generator().manual_seed(8)
    x = torch.randn(30, 41, generator=g)
    x[:, 3] = x[:, 7]                                  # ties
    with _native.emulation(seed=2):
        ex = C.correlation(x, exact=True)
        v = torch.randn(500, generator=g)
        g1, d1 = C.kde(v)
    ref = sps.spearmanr(x.numpy().T).correlation
    assert np.abs(ex.numpy() - ref).max() < 2e-5
    g2, d2 = C.kde(v)
    assert (d1 - d2).abs().max() < 1e-4


def test_replicate_weight_kernel_properties():
    from consensusml_b200.ops.replicate import replicate_weights
    n = 37
    with _native.emulation(seed=1):
        boot = replicate_weights(9, n, mode="bootstrap", seed=13, n_pad=64)
        boot2 = replicate_weights(4, n, mode="bootstrap", seed=13, n_pad=64, rep0=5)
        sub = replicate_weights(9, n, mode="subsample", k=20, seed=13)
        kf = replicate_weights(10, n, mode="kfold", k=5, seed=13)
    assert (boot[:, :n].sum(1) == n).all() and (boot[:, n:] == 0).all() and (boot == boot.round()).all()
    assert torch.equal(boot2, boot[5:9])                      # rep0 continues the same counter-based stream
    assert len({tuple(r.tolist()) for r in boot}) == 9         # replicates differ
    assert ((sub == 0) | (sub == 1)).all() and (sub.sum(1) == 20).all()
    for r in range(2):                                         # every sample is left out by exactly one fold of a repeat
        assert ((kf[5 * r: 5 * r + 5] == 0).sum(0) == 1).all()


# ---- kernels added in round 2: the backend-agnostic cases of tests/kernel_cases.py on the host execution model
import kernel_cases as KC


@pytest.mark.parametrize("case", [n for n in dir(KC) if n.startswith("case_")])
def test_round2_kernel_on_the_host_execution_model(case):
    getattr(KC, case)(KC.EmuBackend())


def test_path_kernels_reproduce_the_stored_glmnet_path_on_the_real_cohort(aml, glmnet_oracle):
    """The reference's own training matrix (93 x 1984) and lambda grid through BOTH path kernels on the host execution
    model, against the coefficient path stored in the reference's lasso_resultslist.rda: same 999 passes as glmnet, the same
    support at lambda.min, coefficients to 1e-5 (fp32 kernel vs float64 Fortran)."""
    import os
    tr, _ = aml.split()
    x, y = tr.design(torch.float32), tr.labels()
    w = torch.ones(2, x.shape[0])
    w[1, 0] = 0                                                       # a leave-one-out fold next to the full fit
    lam = glmnet_oracle["rep1_lambda"]
    ref = np.zeros((1984, 100))
    ref[glmnet_oracle["rep1_train_rows"]] = glmnet_oracle["rep1_train_beta"]
    ref = ref.T
    old = os.environ.get("CML_LASSO_TEAM")
    out = {}
    try:
        for mode in ("0", "1"):
            os.environ["CML_LASSO_TEAM"] = mode
            with _native.emulation(sms=2):
                out[mode] = K.lognet_path(x, y, w, lam)
    finally:
        if old is None:
            os.environ.pop("CML_LASSO_TEAM", None)
        else:
            os.environ["CML_LASSO_TEAM"] = old
    for mode, pb in out.items():
        beta = pb.dense(0).double().numpy()
        assert int(pb.jerr.abs().max()) == 0 and int(pb.npasses[0]) == 999, (mode, pb.npasses)
        assert np.abs(beta - ref).max() < 1e-5
        assert set(np.flatnonzero(beta[89])) == set(np.flatnonzero(ref[89]))
        assert np.abs(pb.dev_ratio[0].double().numpy() - glmnet_oracle["rep1_train_dev_ratio"]).max() < 1e-5
    a, b = out["0"], out["1"]
    assert torch.equal(a.npasses, b.npasses) and torch.equal(a.eta, b.eta) and torch.equal(a.dense(1), b.dense(1))


@pytest.mark.parametrize("k", [3, 6, 8])
def test_voom_wls_with_wide_designs(k):
    """Designs beyond two columns: k = 3 takes the in-register inverse (KMAX = 4), k = 6 / 8 the lane-distributed Gauss-Jordan
    (KMAX = 8, one column of [M | I] per lane).  Oracle: the float64 host restatement of the same voom + lmFit + contrast."""
    from consensusml_b200.stats import de as DE
    c, g = nb_counts(400, 40, seed=20 + k, n_de=60)
    rng = np.random.default_rng(k)
    X = np.column_stack([np.ones(40), np.asarray(g, dtype=np.float64)] + [rng.normal(size=40) for _ in range(k - 2)])
    cvec = np.zeros(k)
    cvec[1] = 1.0
    cvec[k - 1] = -0.5
    ref = DE.voom_de(torch.from_numpy(c), g, design=X, contrast=cvec)                  # CPU: oracle path
    with _native.emulation(seed=2):
        res = DE.voom_de(torch.from_numpy(c), g, design=X, contrast=cvec)
    assert np.array_equal(np.asarray(res.keep), np.asarray(ref.keep))
    assert np.abs(np.asarray(res.logFC) - np.asarray(ref.logFC)).max() < 2e-3
    t_a, t_b = np.asarray(res.t), np.asarray(ref.t)
    assert (np.abs(t_a - t_b) / np.maximum(1.0, np.abs(t_b))).max() < 1e-2
    assert abs(res.s2_prior - ref.s2_prior) / ref.s2_prior < 5e-3


def test_log_cpm_beyond_the_grid_y_limit():
    """More rows than gridDim.y can hold (65,535): the kernel wraps around instead of failing to launch (a transcript-level
    matrix has several hundred thousand rows; the host model enforces the device's launch limits)."""
    from consensusml_b200.stats import normalize as N
    rng = np.random.default_rng(3)
    c = torch.from_numpy(rng.poisson(5, (70001, 3)).astype(np.int32))
    lib = c.sum(0).double()
    ref = N.log_cpm(c, lib)
    with _native.emulation():
        got = N.log_cpm(c, lib)
        got_t = N.log_cpm(c, lib, sample_major=True)
    assert got.shape == (70001, 3) and (got - ref).abs().max() < 1e-5 and torch.equal(got_t.t().contiguous(), got)


def test_latency_kernel_canary_logic(monkeypatch):
    """ops.lasso runs a small batch through both path kernels once per device before latency mode is relied on; here on the host
    model: equal -> 'ok' and no environment left behind; a forced disagreement -> latency mode switched off with a warning."""
    import os
    import warnings
    from consensusml_b200.ops import lasso as K
    monkeypatch.delenv("CML_LASSO_TEAM", raising=False)
    K._CANARY_DONE.clear()
    with _native.emulation(seed=2):
        K._latency_kernel_canary(torch.device("cpu"))
        assert K._CANARY_DONE == {"cpu": "ok"} and "CML_LASSO_TEAM" not in os.environ
        K._CANARY_DONE.clear()
        real = K._path_cuda
        calls = []

        def skewed(*a, **k):
            pb = real(*a, **k)
            calls.append(os.environ.get("CML_LASSO_TEAM"))
            if calls[-1] == "1":
                pb.npasses += 1
            return pb
        monkeypatch.setattr(K, "_path_cuda", skewed)
        with warnings.catch_warnings(record=True) as rec:
            warnings.simplefilter("always")
            K._latency_kernel_canary(torch.device("cpu"))
        assert calls == ["0", "1"] and K._CANARY_DONE == {"cpu": "disabled"} and os.environ.get("CML_LASSO_TEAM") == "0"
        assert any("latency-mode lasso kernel disagreed" in str(w.message) for w in rec)
    monkeypatch.delenv("CML_LASSO_TEAM", raising=False)
    K._CANARY_DONE.clear()


def test_metrics_kernel_first_use_check(monkeypatch):
    """The first answer of cml_binary_metrics on a device is held against the torch restatement; a disagreement makes the
    restatement the implementation for the rest of the process."""
    import warnings
    from consensusml_b200.ops import metrics as M
    g = torch.Generator().manual_seed(2)
    s = torch.randn(3, 50, generator=g)
    t = (torch.rand(50, generator=g) < 0.5).float()
    monkeypatch.setattr(M, "_FORCE_CHECK", True)
    monkeypatch.setattr(M, "_KERNEL_STATE", {})
    with _native.emulation(seed=1):
        _native.reset_calls()
        a = M.batch_metrics(s, t)
        assert M._KERNEL_STATE == {"checked": True} and _native.calls.get("cml_binary_metrics") == 1
        monkeypatch.setattr(M, "_KERNEL_STATE", {})
        real = M._batch_metrics_torch
        monkeypatch.setattr(M, "_batch_metrics_torch", lambda *a_, **k: real(*a_, **k) + torch.tensor([0.01, 0, 0, 0, 0, 0, 0, 0], dtype=torch.float64))
        with warnings.catch_warnings(record=True) as rec:
            warnings.simplefilter("always")
            b = M.batch_metrics(s, t)
        assert M._KERNEL_STATE.get("use") is False and any("disagreed" in str(w.message) for w in rec)
        assert torch.allclose(b["auc"], a["auc"] + 0.01)
        c = M.batch_metrics(s, t)                               # from now on: the restatement, no kernel call
        assert _native.calls.get("cml_binary_metrics") == 2 and torch.allclose(c["auc"], b["auc"])


def test_tree_growth_first_use_check(monkeypatch):
    from consensusml_b200.models import trees as TR
    monkeypatch.setattr(TR, "_FORCE_CANARY", True)
    monkeypatch.setattr(TR, "_CANARY_STATE", {})
    g = torch.Generator().manual_seed(1)
    x = torch.randn(80, 10, generator=g)
    y = (x[:, 0] > 0).float()
    with _native.emulation(seed=2):
        TR.RandomForest(n_trees=4, max_depth=3, seed=1).fit(x, y)
        assert TR._CANARY_STATE == {"cpu": "ok"} and TR.SYNC_FREE_MAX_DEPTH == 8
        # a device on which the two growth modes disagree: the bounded levels are switched off
        monkeypatch.setattr(TR, "_CANARY_STATE", {})
        real = TR.predict_leaves
        n_calls = []

        def skew(tb, bins, m):
            leaf, val = real(tb, bins, m)
            n_calls.append(1)
            return leaf, val + (0.5 if len(n_calls) == 2 else 0.0)
        monkeypatch.setattr(TR, "predict_leaves", skew)
        import warnings
        with warnings.catch_warnings(record=True) as rec:
            warnings.simplefilter("always")
            TR._sync_free_canary(torch.device("cpu"))
        assert TR._CANARY_STATE == {"cpu": "disabled"} and TR.SYNC_FREE_MAX_DEPTH == -1 and any("switched off" in str(w.message) for w in rec)
    monkeypatch.setattr(TR, "SYNC_FREE_MAX_DEPTH", 8)
