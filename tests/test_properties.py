"""Property tests (hypothesis) of invariants the pipeline relies on: order / shard invariance, monotonicity, conservation.
CPU only; sizes are kept small so the whole file runs in seconds."""
import numpy as np
import torch
from hypothesis import given, settings
from hypothesis import strategies as st
from hypothesis.extra import numpy as hnp

from consensusml_b200.consensus import rank_aggregate as RA
from consensusml_b200.ops import colvar as CV
from consensusml_b200.ops import metrics as M
from consensusml_b200.ops.replicate import replicate_weights
from consensusml_b200.oracle import limma as O
from consensusml_b200.parallel.dist import shard_bounds

FAST = settings(max_examples=40, deadline=None)
pvals = hnp.arrays(np.float64, st.integers(2, 60), elements=st.floats(1e-12, 1.0, allow_nan=False))


@FAST
@given(pvals, st.randoms(use_true_random=False))
def test_bh_is_monotone_bounded_and_order_free(p, rnd):
    adj = O.bh_adjust(p)
    assert (adj >= p - 1e-15).all() and (adj <= 1.0).all()
    o = np.argsort(p, kind="stable")
    assert (np.diff(adj[o]) >= -1e-15).all()                       # larger p never gets a smaller adjusted p
    perm = np.array(rnd.sample(range(p.size), p.size))
    assert np.allclose(O.bh_adjust(p[perm]), adj[perm], rtol=0, atol=1e-15)


@FAST
@given(st.integers(1, 200), st.integers(1, 9))
def test_shard_bounds_partition(n, world):
    cuts = [shard_bounds(n, world, r) for r in range(world)]
    assert cuts[0][0] == 0 and cuts[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
    sizes = [hi - lo for lo, hi in cuts]
    assert max(sizes) - min(sizes) <= 1


@FAST
@given(hnp.arrays(np.float64, st.tuples(st.integers(4, 40), st.integers(1, 6)), elements=st.floats(-50, 50, allow_nan=False)),
       st.integers(1, 4))
def test_column_moments_merge_is_shard_count_invariant(x, world):
    t = torch.from_numpy(x)
    full = CV.column_moments(t)
    acc = None
    for r in range(world):
        lo, hi = shard_bounds(t.shape[0], world, r)
        if hi == lo:
            continue
        part = CV.column_moments(t[lo:hi])
        acc = part if acc is None else CV.merge_moments(acc, part)
    for a, b in zip(acc, full):
        assert torch.allclose(torch.as_tensor(a, dtype=torch.float64), torch.as_tensor(b, dtype=torch.float64), rtol=1e-9, atol=1e-9)


@FAST
@given(hnp.arrays(np.float64, st.integers(6, 50), elements=st.floats(-5, 5, allow_nan=False)), st.integers(0, 2 ** 31 - 1))
def test_auc_is_rank_based_and_antisymmetric(score, seed):
    rng = np.random.default_rng(seed)
    y = (rng.random(score.size) < 0.5).astype(np.float64)
    if y.min() == y.max():
        y[0], y[1] = 0.0, 1.0
    s, t = torch.from_numpy(score), torch.from_numpy(y)
    a = M.auc(s, t)
    assert 0.0 <= a <= 1.0
    assert abs(M.auc(s * 4.0, t) - a) < 1e-12                     # order-preserving transform (exact in floating point)
    assert abs(M.auc(-s, t) - (1 - a)) < 1e-12
    perm = torch.from_numpy(rng.permutation(score.size))
    assert abs(M.auc(s[perm], t[perm]) - a) < 1e-12
    cm = M.confusion_matrix((s > 0).long(), t.long())
    assert int(cm.sum()) == score.size


@FAST
@given(hnp.arrays(np.float64, st.tuples(st.integers(5, 40), st.integers(2, 5)), elements=st.floats(-3, 3, allow_nan=False, allow_subnormal=False)),   # a subnormal times 0.5 underflows into a tie
       st.integers(0, 2 ** 31 - 1))
def test_rank_aggregation_is_permutation_equivariant_and_scale_free(imp, seed):
    t = torch.from_numpy(imp)
    ranks = RA.rank_matrix(t)
    assert ranks.min() >= 1 and ranks.max() <= imp.shape[0]
    scaled = RA.rank_matrix(t * torch.tensor([0.5, 7.0, 2.0, 11.0, 3.0])[: imp.shape[1]])
    assert torch.equal(ranks, scaled)                               # per-model positive rescaling changes nothing
    perm = torch.from_numpy(np.random.default_rng(seed).permutation(imp.shape[0]))
    assert torch.allclose(RA.mean_rank(RA.rank_matrix(t[perm])), RA.mean_rank(ranks)[perm])
    rho = RA.robust_rank_aggregation(ranks)
    assert ((rho >= 0) & (rho <= 1 + 1e-12)).all()


@FAST
@given(st.integers(1, 12), st.integers(5, 70), st.integers(0, 2 ** 31 - 1))
def test_replicate_weights_conserve_samples(B, n, seed):
    w = replicate_weights(B, n, seed=seed)[:, :n]
    assert torch.all(w >= 0) and torch.all(w == w.round())
    assert torch.all(w.sum(1) == n)                                  # a bootstrap draws n of n
    k = min(5, n)
    f = replicate_weights(k, n, mode="kfold", k=k, seed=seed)[:, :n]
    held = (f == 0).sum(0)
    assert torch.all(held == 1)                                     # every sample is left out by exactly one fold
    s = replicate_weights(B, n, mode="subsample", k=max(2, n // 2), seed=seed)[:, :n]
    assert torch.all(s.sum(1) == max(2, n // 2)) and s.max() <= 1


@settings(max_examples=8, deadline=None)
@given(st.integers(0, 2 ** 31 - 1))
def test_tmm_factors_have_unit_geometric_mean_and_ignore_sample_order(seed):
    from consensusml_b200.io.synthetic import nb_counts
    c, _ = nb_counts(600, 10, seed=seed % 1000, n_de=60)
    c = c.astype(np.float64)
    lib = c.sum(0)
    f = O.tmm_factors(c, lib)
    assert abs(np.exp(np.log(f).mean()) - 1) < 1e-9
    perm = np.random.default_rng(seed).permutation(10)
    f2 = O.tmm_factors(c[:, perm], lib[perm])
    assert np.allclose(f2, f[perm], rtol=1e-9)


@settings(max_examples=6, deadline=None)
@given(st.integers(0, 2 ** 31 - 1))
def test_lasso_path_does_not_depend_on_sample_order(seed):
    from consensusml_b200.io.synthetic import expression_cohort
    from consensusml_b200.ops import lasso as K
    x, y = expression_cohort(40, 60, n_signal=6, seed=seed % 97, effect=(0.8, 1.5))
    x, y = torch.as_tensor(x).double(), torch.as_tensor(y).double()
    lam = 10.0 ** np.linspace(0, -2, 12)
    a = K.lognet_path(x, y, torch.ones(1, 40, dtype=torch.float64), lam, thresh=1e-10)
    perm = torch.from_numpy(np.random.default_rng(seed).permutation(40))
    b = K.lognet_path(x[perm], y[perm], torch.ones(1, 40, dtype=torch.float64), lam, thresh=1e-10)
    assert (a.dense(0) - b.dense(0)).abs().max() < 1e-6
