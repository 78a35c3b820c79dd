"""Random forests and gradient-boosted trees on the batched histogram tree engine (csrc/trees.cu).

* :class:`RandomForest`    - Breiman RF as used at cml_targetaml_seanalysis.Rmd:1037-1050 (n-of-n bootstrap,
  ``mtry = floor(sqrt(p))``, ``nodesize = 1``, Gini splits, OOB error, MeanDecreaseGini, proximity) and the
  depth-limited sklearn variant of scripts/model_comp.py:9.
* :class:`GradientBoosting` - second-order (xgboost: model_comp.py:7, cml_targetaml_seanalysis.Rmd:1164-1186)
  and first-order stochastic (sklearn ``GradientBoostingClassifier``: model_comp.py:8) boosting with logistic
  loss and gain / cover / frequency importance.

Every tree of a batch grows in the same launches (one per level).  With n <= 256 samples features are
rank-binned, so the split search equals the exact greedy search of the CPU libraries.  CPU tensors run a
NumPy restatement of the same level-wise algorithm (``_grow_numpy``), which is also the test oracle.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np
import torch

from .. import _native

_I, _F, _P, _ULL = ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_ulonglong
_native.register_signature("cml_rank_bins", [_P, _I, _I, _I, _P, _I, _P, _P, _P])
_native.register_signature("cml_apply_bins", [_P, _I, _I, _I, _P, _P, _P, _I, _P])
_native.register_signature("cml_forest_level", [_P, _I, _I, _I, _I, _P, _P, _P, _I, _I, _I, _P, _P, _P, _I, _P, _P, _P,
                                                _P, _P, _P, _P, _P, _P, _I, _I, _F, _F, _F, _F, _F, _F, _ULL, _P, _P, _P,
                                                _P, _P])
# (n_frontier_dev, frontier_bound) first; then cml_forest_level's arguments without its host-side n_frontier
_native.register_signature("cml_forest_level_dev", [_P, _I, _P, _I, _I, _I, _I, _P, _P, _P, _I, _I, _I, _P, _P, _P, _P, _P, _P,
                                                    _P, _P, _P, _P, _P, _P, _I, _I, _F, _F, _F, _F, _F, _F, _ULL, _P, _P, _P,
                                                    _P, _P])
_native.register_signature("cml_forest_predict", [_P, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P])
_native.register_signature("cml_forest_proximity", [_P, _P, _I, _I, _P, _P, _P])
_native.register_signature("cml_forest_onehot", [_P, _P, _I, _I, _I, _P, ctypes.c_longlong, _P, ctypes.c_longlong, _P])
_native.register_signature("cml_quantile_edges", [_P, _I, _I, ctypes.c_longlong, _I, _P, _P, _P])

GINI, NEWTON = 0, 1
QUANTILE_SAMPLE = 16384       # values per gene that enter the quantile sort (csrc/select.cu kQMaxSample)


# ------------------------------------------------------------------------------------------------ binning
@dataclass
class Binner:
    """Per-feature bins.  ``edges[f, b]`` is the value bin ``b`` stands for (distinct values when n <= 256,
    quantile representatives otherwise); new data are binned at the midpoints between neighbours."""
    edges: torch.Tensor      # [p, 256] fp32
    n_edges: torch.Tensor    # [p] int32
    max_bin: int

    @staticmethod
    def fit(x: torch.Tensor, max_bin: int = 256):
        """x: samples x genes.  Returns (binner, bins[p, n_pad] uint8)."""
        n, p = x.shape
        n_pad = (n + 3) // 4 * 4
        xt = x.t().contiguous().float()
        if n <= 256 and _native.on_device(x):
            bins = torch.empty(p, n_pad, dtype=torch.uint8, device=x.device)
            edges = torch.full((p, 256), float("inf"), dtype=torch.float32, device=x.device)
            n_edges = torch.empty(p, dtype=torch.int32, device=x.device)
            rc = _native.cuda().cml_rank_bins(_native.ptr(xt), p, n, n, _native.ptr(bins), n_pad, _native.ptr(edges),
                                              _native.ptr(n_edges), _native.stream_ptr())
            _native.check(rc, "cml_rank_bins")
            return Binner(edges, n_edges, int(n_edges.max().item())), bins
        if n > 256 and _native.on_device(x):
            # quantile representatives from a shared-memory sort per gene (csrc/select.cu), then the midpoint binning kernel
            edges = torch.empty(p, 256, dtype=torch.float32, device=x.device)
            n_edges = torch.empty(p, dtype=torch.int32, device=x.device)
            rc = _native.cuda().cml_quantile_edges(_native.ptr(xt), p, n, n, int(max_bin), _native.ptr(edges), _native.ptr(n_edges),
                                                   _native.stream_ptr())
            _native.check(rc, "cml_quantile_edges")
            b = Binner(edges, n_edges, max(2, int(n_edges.max().item())))
            return b, b.transform(x)
        xs = xt.double()
        if n <= 256:
            reps = [torch.unique(xs[f]) for f in range(p)]
        else:
            # same rule as the kernel: all values up to 16,384 samples, beyond that an evenly strided sample of 16,384;
            # type-7 quantiles in float64, rounded to fp32, duplicates dropped
            S = min(n, QUANTILE_SAMPLE)
            sub = xs if S == n else xs[:, (torch.arange(S, dtype=torch.int64, device=x.device) * n) // S]
            qs = torch.linspace(0, 1, max_bin, dtype=torch.float64, device=x.device)
            qv = torch.quantile(sub, qs, dim=1).t().float()
            reps = [torch.unique_consecutive(qv[f]) for f in range(p)]
        edges = torch.full((p, 256), float("inf"), dtype=torch.float32, device=x.device)
        n_edges = torch.zeros(p, dtype=torch.int32, device=x.device)
        for f, r in enumerate(reps):
            edges[f, : r.numel()] = r.float()
            n_edges[f] = r.numel()
        b = Binner(edges, n_edges, max(2, int(n_edges.max().item())))
        return b, b.transform(x)

    def transform(self, x: torch.Tensor) -> torch.Tensor:
        m, p = x.shape
        m_pad = (m + 3) // 4 * 4
        xt = x.t().contiguous().float()
        if _native.on_device(x):
            bins = torch.empty(p, m_pad, dtype=torch.uint8, device=x.device)
            rc = _native.cuda().cml_apply_bins(_native.ptr(xt), p, m, m, _native.ptr(self.edges), _native.ptr(self.n_edges),
                                               _native.ptr(bins), m_pad, _native.stream_ptr())
            _native.check(rc, "cml_apply_bins")
            return bins
        bins = torch.zeros(p, m_pad, dtype=torch.uint8)
        e, ne, xn, out = self.edges.numpy(), self.n_edges.numpy(), xt.numpy(), bins.numpy()
        for f in range(p):
            mids = 0.5 * (e[f, : ne[f] - 1] + e[f, 1: ne[f]])
            out[f, :m] = np.searchsorted(mids, xn[f], side="right")
        return bins


# ------------------------------------------------------------------------------------------------ tree batch
@dataclass
class TreeBatch:
    feat: torch.Tensor       # [T, max_nodes] int32, -1 = leaf
    thr: torch.Tensor        # [T, max_nodes] int32 (go left iff bin <= thr)
    left: torch.Tensor       # [T, max_nodes] int32 (right child = left + 1)
    value: torch.Tensor      # [T, max_nodes] fp32
    gain: torch.Tensor
    cover: torch.Tensor
    node_count: torch.Tensor # [T] int32


def grow(bins: torch.Tensor, n: int, s0: torch.Tensor, s1: torch.Tensor, *, s2: Optional[torch.Tensor] = None,
         criterion: int = GINI, mtry: int = 0, max_depth: int = 64, max_nodes: Optional[int] = None, max_bin: int = 256,
         lam: float = 0.0, gamma: float = 0.0, min_child_weight: float = 0.0, min_samples_leaf: float = 1.0,
         min_samples_split: float = 2.0, eta: float = 1.0, seed: int = 0, feature_mask: Optional[torch.Tensor] = None,
         importance: Optional[Dict[str, torch.Tensor]] = None, in_tree: Optional[torch.Tensor] = None) -> TreeBatch:
    """Grow ``T = s0.shape[0]`` trees.  ``in_tree`` ([T, n] bool) restricts each tree to a sample subset
    (bootstrap members / row subsample); samples outside never enter the root."""
    T = s0.shape[0]
    p, n_pad = bins.shape
    dev = bins.device
    if max_nodes is None:
        max_nodes = min(2 * n + 1, 2 ** (min(max_depth, 20) + 1) - 1)
    if not _native.on_device(bins):
        return _grow_numpy(bins, n, s0, s1, s2, criterion, mtry, max_depth, max_nodes, max_bin, lam, gamma, min_child_weight,
                           min_samples_leaf, min_samples_split, eta, seed, feature_mask, importance, in_tree)
    if bins.is_cuda or _FORCE_CANARY:
        _sync_free_canary(dev)
    i32 = dict(dtype=torch.int32, device=dev)
    f32 = dict(dtype=torch.float32, device=dev)
    if in_tree is None:
        in_tree = torch.ones(T, n, dtype=torch.bool, device=dev)
    order = torch.argsort((~in_tree).to(torch.uint8), dim=1, stable=True).to(torch.int32).contiguous()
    counts = in_tree.sum(1).to(torch.int32)
    order2 = order.clone()
    feat = torch.full((T, max_nodes), -1, **i32)
    thr = torch.zeros(T, max_nodes, **i32)
    left = torch.full((T, max_nodes), -1, **i32)
    value = torch.zeros(T, max_nodes, **f32)
    gain = torch.zeros(T, max_nodes, **f32)
    cover = torch.zeros(T, max_nodes, **f32)
    node_count = torch.ones(T, **i32)
    frontier = torch.zeros(T, 7, **i32)
    frontier[:, 0] = torch.arange(T, **i32)
    frontier[:, 3] = counts
    nf = T
    next_count = torch.zeros(1, **i32)
    lb = _native.cuda()
    P = _native.ptr
    s0 = s0.float().contiguous()
    s1 = s1.float().contiguous()
    s2c = s2.float().contiguous() if s2 is not None else None
    fm = feature_mask.to(device=dev, dtype=torch.uint8).contiguous() if feature_mask is not None else None
    imp = importance or {}
    depth = 0
    # The first levels are launched WITHOUT host synchronisation: the frontier of level d has at most T * 2^d entries, so each of
    # them gets that bound as its grid and reads the true size on the device (the previous level's counter); no counter reset, no
    # frontier compaction.  Shallow trees (boosting rounds of depth 2 - 8) finish here; deeper ones pay their first host round trip
    # only at the level where the bound would exceed a modest grid, and continue with the synchronised, compacted loop below.
    free_levels = 0
    while free_levels <= min(max_depth, SYNC_FREE_MAX_DEPTH) and T * (2 ** free_levels) <= SYNC_FREE_MAX_BLOCKS:
        free_levels += 1
    if free_levels > 0:
        counts = torch.zeros(free_levels + 1, **i32)
        for depth in range(free_levels):
            bound = min(T * (2 ** depth), T * max_nodes)
            nxt = torch.empty(2 * bound, 7, **i32)
            rc = lb.cml_forest_level_dev(P(counts[depth:]) if depth > 0 else None, bound, P(bins), p, n, n_pad, max_bin, P(s0), P(s1), P(s2c), T,
                                         max_nodes, max_depth, P(order), P(order2), P(frontier), P(nxt), P(counts[depth + 1:]), P(node_count),
                                         P(feat), P(thr), P(left), P(value), P(gain), P(cover), criterion, int(mtry), float(lam), float(gamma),
                                         float(min_child_weight), float(min_samples_leaf), float(min_samples_split), float(eta),
                                         seed & (2**64 - 1), P(fm), P(imp.get("gain")), P(imp.get("cover")), P(imp.get("freq")),
                                         _native.stream_ptr())
            _native.check(rc, "cml_forest_level_dev")
            frontier = nxt
            order, order2 = order2, order
        depth = free_levels
        if depth > max_depth:
            return TreeBatch(feat, thr, left, value, gain, cover, node_count)
        nf = int(counts[free_levels].item())
        frontier = frontier[:nf].contiguous()
    while nf > 0 and depth <= max_depth:
        nxt = torch.empty(2 * nf, 7, **i32)
        next_count.zero_()
        rc = lb.cml_forest_level(P(bins), p, n, n_pad, max_bin, P(s0), P(s1), P(s2c), T, max_nodes, max_depth, P(order),
                                 P(order2), P(frontier), nf, P(nxt), P(next_count), P(node_count), P(feat), P(thr), P(left),
                                 P(value), P(gain), P(cover), criterion, int(mtry), float(lam), float(gamma),
                                 float(min_child_weight), float(min_samples_leaf), float(min_samples_split), float(eta),
                                 seed & (2**64 - 1), P(fm), P(imp.get("gain")), P(imp.get("cover")), P(imp.get("freq")),
                                 _native.stream_ptr())
        _native.check(rc, "cml_forest_level")
        nf = int(next_count.item())
        frontier = nxt[:nf].contiguous()
        order, order2 = order2, order
        depth += 1
    return TreeBatch(feat, thr, left, value, gain, cover, node_count)


_CANARY_STATE: dict = {}
_FORCE_CANARY = False           # tests: run the first-use comparison on the host execution model too


def _sync_free_canary(dev) -> None:
    """Once per process and device, before the first real growth: a small forest grown with the bounded, synchronisation-free
    levels and with the synchronised loop.  The former has run on the host execution model only; if the leaf values the two
    assign to the training samples differ on this device, it is switched off for the process (loudly)."""
    global SYNC_FREE_MAX_DEPTH
    key = str(dev)
    if key in _CANARY_STATE or SYNC_FREE_MAX_DEPTH < 0:
        return
    _CANARY_STATE[key] = "running"                       # grow() below re-enters this function
    g = torch.Generator().manual_seed(5)
    n, p, T = 64, 12, 6
    xb = torch.randint(0, 16, (p, n), generator=g, dtype=torch.uint8).to(dev)
    yv = (torch.rand(T, n, generator=g) < 0.5).float().to(dev)
    w = torch.ones(T, n, device=dev)
    outs = []
    keep = SYNC_FREE_MAX_DEPTH
    try:
        for mode in (keep, -1):
            SYNC_FREE_MAX_DEPTH = mode
            tb = grow(xb, n, w * yv, w, criterion=GINI, mtry=4, max_depth=5, max_bin=16, seed=3)
            outs.append(predict_leaves(tb, xb, n)[1].cpu())
    finally:
        SYNC_FREE_MAX_DEPTH = keep
    if torch.allclose(outs[0], outs[1], atol=1e-6):
        _CANARY_STATE[key] = "ok"
    else:
        import warnings
        warnings.warn("consensusml_b200: tree growth without per-level synchronisation disagreed with the synchronised loop on this "
                      "device; it is switched off for this process")
        SYNC_FREE_MAX_DEPTH = -1
        _CANARY_STATE[key] = "disabled"


def predict_leaves(trees: TreeBatch, bins: torch.Tensor, m: int):
    """(leaf[T, m] int32, value[T, m] fp32) for samples binned as ``bins[p, m_pad]``."""
    T, max_nodes = trees.feat.shape
    if not _native.on_device(bins):
        return _predict_numpy(trees, bins, m)
    leaf = torch.empty(T, m, dtype=torch.int32, device=bins.device)
    val = torch.empty(T, m, dtype=torch.float32, device=bins.device)
    P = _native.ptr
    rc = _native.cuda().cml_forest_predict(P(bins), m, bins.shape[1], T, max_nodes, P(trees.feat), P(trees.thr), P(trees.left),
                                           P(trees.value), P(leaf), P(val), _native.stream_ptr())
    _native.check(rc, "cml_forest_predict")
    return leaf, val


SYNC_FREE_MAX_DEPTH = 8         # grow(): trees this shallow are grown without a host synchronisation per level ...
SYNC_FREE_MAX_BLOCKS = 1 << 15  # ... as long as the widest level's upper bound stays a modest grid
PROXIMITY_GEMM_MIN_N = 256      # below this the O(n^2 T) comparison kernel is one cheap launch; above it the GEMM form wins


def proximity_gemm(leaf: torch.Tensor, oob: Optional[torch.Tensor], max_nodes: int, max_cols: int = 1 << 16):
    """(proximity counts, pair counts) [n, n] of a batch of trees as one-hot GEMMs: ``forest_onehot_kernel`` writes the
    sample x (tree, leaf) membership matrix in bf16 and the tcgen05 GEMM multiplies it with its transpose (randomForest's
    ``proximity = TRUE``, cml_targetaml_seanalysis.Rmd:1037-1050, at cohort sizes where n^2 T comparisons are too many).
    Trees are processed in groups of at most ``max_cols`` one-hot columns."""
    from ..ops import gemm as G
    T, n = leaf.shape
    dev = leaf.device
    per = max(1, max_cols // max_nodes)
    prox = torch.zeros(n, n, device=dev)
    pairs = torch.zeros(n, n, device=dev)
    for t0 in range(0, T, per):
        t1 = min(T, t0 + per)
        tt = t1 - t0
        ldz = (tt * max_nodes + 7) // 8 * 8
        ldo = (tt + 7) // 8 * 8
        z = torch.zeros(n, ldz, dtype=torch.bfloat16, device=dev)
        o = torch.zeros(n, ldo, dtype=torch.bfloat16, device=dev)
        lf = leaf[t0:t1].contiguous()
        ob = oob[t0:t1].contiguous() if oob is not None else None
        rc = _native.cuda().cml_forest_onehot(_native.ptr(lf), _native.ptr(ob), tt, n, max_nodes, _native.ptr(z), ldz, _native.ptr(o), ldo,
                                              _native.stream_ptr())
        _native.check(rc, "cml_forest_onehot")
        # tcgen05 GEMMs (on the host execution model: the same kernels on the model of the asynchronous units)
        prox += G.gemm_tn(z, [z], n) if ldz >= G.MIN_K else z.float() @ z.float().t()
        pairs += G.gemm_tn(o, [o], n) if ldo >= G.MIN_K else o.float() @ o.float().t()
    return prox, pairs


# ------------------------------------------------------------------------------------------------ NumPy restatement
def _score(crit, a, b, lam):
    if crit == GINI:
        return np.where(b > 0, (a * a + (b - a) * (b - a)) / np.where(b > 0, b, 1), 0.0)
    den = b + lam
    return np.where(den > 0, a * a / np.where(den > 0, den, 1), 0.0)


_M64 = (1 << 64) - 1


def _mix(z):
    z = (z + 0x9E3779B97F4A7C15) & _M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def _node_candidates(seed, tree, key, p, mtry, fmask):
    """Same counter-based sampling as the kernel (splitmix64 keyed by the node's path from the root), so CPU
    and GPU trees see the same candidates whatever order node ids were handed out in."""
    base = _mix((seed & _M64) ^ ((0x9E3779B97F4A7C15 * (tree + 1)) & _M64) ^ key)
    got, seen, ctr = [], set(), 0
    while len(got) < mtry and ctr < 64 * p:
        f = ((_mix((base + ctr) & _M64) >> 32) * p) >> 32
        ctr += 1
        if f in seen or (fmask is not None and not fmask[f]):
            continue
        seen.add(f)
        got.append(f)
    return got


def _grow_numpy(bins, n, s0, s1, s2, crit, mtry, max_depth, max_nodes, max_bin, lam, gamma, mcw, msl, mss, eta, seed, fmask,
                importance, in_tree):
    B = bins.numpy()
    S0, S1 = s0.double().numpy(), s1.double().numpy()
    S2 = s2.double().numpy() if s2 is not None else None
    T, p = S0.shape[0], B.shape[0]
    fm = fmask.numpy().astype(bool) if fmask is not None else None
    feat = np.full((T, max_nodes), -1, np.int32)
    thr = np.zeros((T, max_nodes), np.int32)
    left = np.full((T, max_nodes), -1, np.int32)
    value = np.zeros((T, max_nodes), np.float32)
    gain = np.zeros((T, max_nodes), np.float32)
    cover = np.zeros((T, max_nodes), np.float32)
    ncount = np.ones(T, np.int32)
    all_feats = mtry <= 0 or mtry >= p
    for t in range(T):
        members = np.arange(n) if in_tree is None else np.flatnonzero(in_tree[t].numpy())
        frontier = [(0, members, 0, 0)]
        while frontier:
            nxt = []
            for node, idx, depth, key in frontier:
                a, b = S0[t, idx].sum(), S1[t, idx].sum()
                if crit == GINI:
                    v = a / b if b > 0 else 0.0
                elif S2 is not None:
                    c = S2[t, idx].sum()
                    v = eta * (a / c if c > 1e-30 else 0.0)
                else:
                    v = -eta * a / (b + lam)
                value[t, node], cover[t, node] = v, b
                pure = crit == GINI and (a <= 0 or a >= b)
                if not (depth < max_depth and idx.size >= mss and idx.size >= 2 and not pure and b >= 2 * mcw):
                    continue
                cands = [f for f in range(p) if fm is None or fm[f]] if all_feats else _node_candidates(seed, t, key, p, mtry, fm)
                parent = float(_score(crit, np.float64(a), np.float64(b), lam))
                best = (-np.inf, -1, -1)
                for f in cands:
                    bb = B[f, idx]
                    h0 = np.bincount(bb, weights=S0[t, idx], minlength=max_bin)
                    h1 = np.bincount(bb, weights=S1[t, idx], minlength=max_bin)
                    c0, c1 = np.cumsum(h0), np.cumsum(h1)
                    r0, r1 = a - c0, b - c1
                    ok = (h1 > 0) & (r1 > 0) & (c1 >= mcw) & (r1 >= mcw) & (c1 >= msl) & (r1 >= msl)
                    ok[max_bin - 1:] = False
                    if not ok.any():
                        continue
                    g = np.where(ok, _score(crit, c0, c1, lam) + _score(crit, r0, r1, lam) - parent, -np.inf)
                    bi = int(np.argmax(g))
                    if g[bi] > best[0] or (g[bi] == best[0] and (f < best[1] or best[1] < 0)):
                        best = (float(g[bi]), f, bi)
                bg, bf, bbin = best
                real = bg - gamma if crit == NEWTON else bg
                if bf < 0 or not real > 1e-7 * max(1.0, abs(parent)):
                    continue
                child = int(ncount[t])
                if child + 1 >= max_nodes:
                    continue
                ncount[t] += 2
                goes_left = B[bf, idx] <= bbin
                feat[t, node], thr[t, node], left[t, node], gain[t, node] = bf, bbin, child, real
                if importance is not None:
                    importance["gain"][bf] += real
                    importance["cover"][bf] += b
                    importance["freq"][bf] += 1
                nxt.append((child, idx[goes_left], depth + 1, _mix(key ^ 0xA24BAED4963EE407)))
                nxt.append((child + 1, idx[~goes_left], depth + 1, _mix(key ^ 0x9FB21C651E98DF25)))
            frontier = nxt
    t_ = torch.from_numpy
    return TreeBatch(t_(feat), t_(thr), t_(left), t_(value), t_(gain), t_(cover), t_(ncount))


def _predict_numpy(trees: TreeBatch, bins, m):
    B = bins.numpy()
    feat, thr, left, value = (x.numpy() for x in (trees.feat, trees.thr, trees.left, trees.value))
    T = feat.shape[0]
    leaf = np.zeros((T, m), np.int32)
    cols = np.arange(m)
    for t in range(T):
        node = np.zeros(m, np.int64)
        while True:
            f = feat[t, node]
            ia = np.flatnonzero(f >= 0)
            if ia.size == 0:
                break
            go_right = B[f[ia], cols[ia]] > thr[t, node[ia]]
            node[ia] = left[t, node[ia]] + go_right
        leaf[t] = node
    val = np.take_along_axis(value, leaf.astype(np.int64), axis=1)
    return torch.from_numpy(leaf), torch.from_numpy(val.astype(np.float32))


# ------------------------------------------------------------------------------------------------ random forest
@dataclass
class RandomForest:
    """Binary-classification random forest (Breiman)."""
    n_trees: int = 500
    mtry: Optional[int] = None            # default floor(sqrt(p)) like randomForest
    max_depth: int = 64
    min_samples_leaf: int = 1             # randomForest nodesize
    bootstrap: bool = True
    seed: int = 20                        # the reference's set.seed(20), cml_targetaml_seanalysis.Rmd:1026
    proximity: bool = False
    oob_prox: bool = True                 # randomForest's default: proximity from out-of-bag pairs only
    tree_batch: int = 2048

    binner: Binner = field(default=None, repr=False)
    trees: list = field(default_factory=list, repr=False)
    importance_: torch.Tensor = field(default=None, repr=False)      # MeanDecreaseGini
    oob_error_: float = None
    oob_votes_: torch.Tensor = field(default=None, repr=False)
    proximity_: torch.Tensor = field(default=None, repr=False)
    confusion_oob_: torch.Tensor = field(default=None, repr=False)

    def fit(self, x: torch.Tensor, y: torch.Tensor) -> "RandomForest":
        self.begin(x, y)
        parts, done = [], 0
        while done < self.n_trees:
            T = min(self.tree_batch, self.n_trees - done)
            parts.append(self.grow_chunk(done, T))
            done += T
        return self.finish(parts)

    # The forest is a sum over trees of everything it reports (importance, out-of-bag votes, proximity counts), and a
    # tree's randomness depends only on (seed, tree index).  ``begin / grow_chunk / finish`` expose that: chunks of
    # trees can be grown in any order, on any rank (parallel/tasks.py), and merged - the result does not depend on how
    # the chunks were distributed, only on where their boundaries are.
    def begin(self, x: torch.Tensor, y: torch.Tensor) -> "RandomForest":
        n, p = x.shape
        self.binner, self._bins = Binner.fit(x)
        self._mtry = self.mtry if self.mtry is not None else max(1, int(np.floor(np.sqrt(p))))
        self._yf = y.float().to(x.device)
        self._n, self._p = n, p
        self.trees = []
        return self

    def grow_chunk(self, tree0: int, T: int, x_eval: Optional[torch.Tensor] = None, keep_trees: bool = True) -> Dict[str, torch.Tensor]:
        """Grow trees [tree0, tree0 + T) and return their contribution to every forest statistic (device tensors):
        importance sums, out-of-bag vote counts, proximity / pair counts and - with ``x_eval`` - the number of trees
        voting class 1 for each evaluation sample."""
        from ..ops.replicate import replicate_weights
        n, p, dev, bins, yf = self._n, self._p, self._bins.device, self._bins, self._yf
        imp = {k: torch.zeros(p, dtype=torch.float32, device=dev) for k in ("gain", "cover", "freq")}
        if self.bootstrap:
            w = replicate_weights(T, n, mode="bootstrap", seed=self.seed, device=dev, rep0=tree0)[:, :n].contiguous()
        else:
            w = torch.ones(T, n, device=dev)
        tb = grow(bins, n, w * yf[None, :], w, criterion=GINI, mtry=self._mtry, max_depth=self.max_depth,
                  max_bin=self.binner.max_bin, min_samples_leaf=float(self.min_samples_leaf), seed=self.seed + tree0,
                  importance=imp, in_tree=w > 0)
        if keep_trees:
            self.trees.append(tb)
        leaf, val = predict_leaves(tb, bins, n)
        oob = (w == 0).float()
        part = {"n_trees": torch.tensor(T), "gain": imp["gain"], "cover": imp["cover"], "freq": imp["freq"],
                "votes1": ((val > 0.5).float() * oob).sum(0), "nvote": oob.sum(0), "nodes": tb.node_count.sum()}
        if self.proximity:
            prox = torch.zeros(n, n, device=dev)
            pairs = torch.zeros(n, n, device=dev)
            use_oob = self.oob_prox and self.bootstrap
            ob = (w == 0).to(torch.uint8).contiguous() if use_oob else None
            if _native.on_device(leaf) and n >= PROXIMITY_GEMM_MIN_N:
                dp, dq = proximity_gemm(leaf.contiguous(), ob, tb.feat.shape[1])
                prox += dp
                pairs += dq
            elif _native.on_device(leaf):
                rc = _native.cuda().cml_forest_proximity(_native.ptr(leaf.contiguous()), _native.ptr(ob), T, n,
                                                         _native.ptr(prox), _native.ptr(pairs), _native.stream_ptr())
                _native.check(rc, "cml_forest_proximity")
            else:
                both = (ob[:, :, None] & ob[:, None, :]).bool() if use_oob else torch.ones(T, n, n, dtype=torch.bool)
                prox += ((leaf[:, :, None] == leaf[:, None, :]) & both).float().sum(0)
                pairs += both.float().sum(0)
            part["prox"], part["pairs"] = prox, pairs
        if x_eval is not None:
            _, ve = predict_leaves(tb, self.binner.transform(x_eval), x_eval.shape[0])
            part["eval_votes1"] = (ve > 0.5).float().sum(0)
        return part

    def finish(self, parts) -> "RandomForest":
        """Merge chunk contributions (in the order given; sums of non-negative counts, so the order does not matter
        for anything but the last bit of the fp32 importance sums - pass them sorted by tree0 for bit-reproducibility)."""
        from ..ops.metrics import confusion_matrix
        dev = self._bins.device
        tot = {}
        for part in parts:
            for k, v in part.items():
                v = torch.as_tensor(v).to(dev)
                tot[k] = v.clone() if k not in tot else tot[k] + v
        nt = int(tot["n_trees"])
        self.n_trees_grown_ = nt
        self.importance_ = tot["gain"] / nt
        self.importance_cover_ = tot["cover"]
        self.importance_freq_ = tot["freq"]
        self.mean_nodes_ = float(tot["nodes"]) / nt
        frac = tot["votes1"] / tot["nvote"].clamp_min(1)
        self.oob_votes_ = frac
        pred = (frac > 0.5).float()
        has = tot["nvote"] > 0
        yf = self._yf
        self.oob_error_ = float((pred[has] != yf[has]).float().mean())
        self.confusion_oob_ = confusion_matrix(pred[has], yf[has])
        if "prox" in tot:
            self.proximity_ = tot["prox"] / tot["pairs"].clamp_min(1)
            self.proximity_.fill_diagonal_(1.0)
        self.eval_proba_ = tot["eval_votes1"] / nt if "eval_votes1" in tot else None
        return self

    def _accumulate(self, x, fn):
        bins = self.binner.transform(x)
        m = x.shape[0]
        acc = torch.zeros(m, device=x.device)
        for tb in self.trees:
            _, val = predict_leaves(tb, bins, m)
            acc += fn(val).sum(0)
        return acc / self.n_trees

    def predict_proba(self, x: torch.Tensor) -> torch.Tensor:
        """Fraction of trees voting class 1 (randomForest ``type="prob"``)."""
        return self._accumulate(x, lambda v: (v > 0.5).float())

    def predict_proba_mean(self, x: torch.Tensor) -> torch.Tensor:
        """Mean of per-tree leaf class fractions (sklearn ``predict_proba``)."""
        return self._accumulate(x, lambda v: v)

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return (self.predict_proba(x) > 0.5).long()


# ------------------------------------------------------------------------------------------------ gradient boosting
@dataclass
class GradientBoosting:
    """Logistic-loss tree boosting.  ``mode='xgboost'``: second order with L2 ``reg_lambda`` (XGBClassifier as
    used in model_comp.py:7).  ``mode='sklearn'``: first-order residual fitting with friedman-MSE splits, Newton
    leaf values, row ``subsample`` and per-node ``max_features`` (GradientBoostingClassifier, model_comp.py:8)."""
    n_estimators: int = 100
    learning_rate: float = 0.3
    max_depth: int = 6
    reg_lambda: float = 1.0
    gamma: float = 0.0
    min_child_weight: float = 1.0
    subsample: float = 1.0
    max_features: Optional[object] = None     # None | 'log2' | 'sqrt' | int
    min_samples_leaf: int = 1
    mode: str = "xgboost"
    seed: int = 8
    base_score: float = 0.5

    binner: Binner = field(default=None, repr=False)
    trees: list = field(default_factory=list, repr=False)
    train_error_: list = field(default_factory=list, repr=False)
    importance_: Dict[str, torch.Tensor] = field(default=None, repr=False)
    init_: float = 0.0

    def _mtry(self, p):
        mf = self.max_features
        if mf is None:
            return 0
        if mf == "log2":
            return max(1, int(np.log2(p)))
        if mf == "sqrt":
            return max(1, int(np.sqrt(p)))
        return int(mf)

    def fit(self, x: torch.Tensor, y: torch.Tensor) -> "GradientBoosting":
        from ..ops.replicate import replicate_weights
        n, p = x.shape
        dev = x.device
        self.binner, bins = Binner.fit(x)
        yf = y.float().to(dev)
        imp = {k: torch.zeros(p, dtype=torch.float32, device=dev) for k in ("gain", "cover", "freq")}
        if self.mode == "xgboost":
            self.init_ = float(np.log(self.base_score / (1 - self.base_score)))
        else:
            pr = float(yf.mean())
            self.init_ = float(np.log(pr / (1 - pr)))
        margin = torch.full((n,), self.init_, device=dev)
        self.trees, self.train_error_ = [], []
        k_sub = max(1, int(round(self.subsample * n)))
        for r in range(self.n_estimators):
            pr = torch.sigmoid(margin)
            in_tree = None
            if self.subsample < 1.0:
                in_tree = replicate_weights(1, n, mode="subsample", k=k_sub, seed=self.seed, device=dev, rep0=r)[:, :n] > 0
            if self.mode == "xgboost":
                g, h = (pr - yf)[None, :], (pr * (1 - pr))[None, :]
                tb = grow(bins, n, g, h, criterion=NEWTON, mtry=self._mtry(p), max_depth=self.max_depth,
                          max_bin=self.binner.max_bin, lam=self.reg_lambda, gamma=self.gamma,
                          min_child_weight=self.min_child_weight, min_samples_leaf=0.0, eta=self.learning_rate,
                          seed=self.seed + r, importance=imp, in_tree=in_tree)
            else:
                res = (yf - pr)[None, :]
                tb = grow(bins, n, res, torch.ones_like(res), s2=(pr * (1 - pr))[None, :], criterion=NEWTON,
                          mtry=self._mtry(p), max_depth=self.max_depth, max_bin=self.binner.max_bin, lam=0.0,
                          min_samples_leaf=float(self.min_samples_leaf), eta=self.learning_rate, seed=self.seed + r,
                          importance=imp, in_tree=in_tree)
            self.trees.append(tb)
            _, val = predict_leaves(tb, bins, n)
            margin = margin + val[0]
            self.train_error_.append(((margin > 0).float() != yf).float().mean())      # stays on the device: no host sync per round
        self.train_error_ = torch.stack(self.train_error_).cpu().tolist() if self.train_error_ else []
        tot = {k: float(v.sum()) for k, v in imp.items()}
        self.importance_ = {"gain": imp["gain"] / max(tot["gain"], 1e-30), "cover": imp["cover"] / max(tot["cover"], 1e-30),
                            "frequency": imp["freq"] / max(tot["freq"], 1e-30)}
        return self

    def decision_function(self, x: torch.Tensor, n_trees: Optional[int] = None) -> torch.Tensor:
        bins = self.binner.transform(x)
        m = x.shape[0]
        margin = torch.full((m,), self.init_, device=x.device)
        for tb in self.trees[: n_trees or len(self.trees)]:
            _, val = predict_leaves(tb, bins, m)
            margin = margin + val.sum(0)
        return margin

    def predict_proba(self, x: torch.Tensor, n_trees: Optional[int] = None) -> torch.Tensor:
        return torch.sigmoid(self.decision_function(x, n_trees))

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return (self.decision_function(x) > 0).long()
