"""Sample-sharded gradient-boosted trees for large cohorts (BASELINE.json config 4: 1M samples x 20k genes).

The small-cohort engine (models/trees.py) grows whole trees inside CTAs; at this size the per-level gradient
histogram is the job.  Every rank streams its uint8-binned sample shard once per level; the (g, h) histograms of the
open nodes are reduced across ranks INSIDE the build kernel and land, feature by feature, on the rank that owns the
feature (``parallel.fused.HistEngine``: shared-memory int32 accumulation -> rank-local tile -> peer ``red.add`` into the
owner's buffer).  The owner scores the splits of its features; one MAX all-reduce of n_nodes packed keys picks the
winner, so the decision is single-sourced by construction.  Only the smaller child of every split is histogrammed
(sibling = parent - child, derived on the owner), gradients, node totals and sample routing are kernels
(csrc/gbm.cu), and a boosting level issues no host synchronisation: it is a fixed sequence of launches plus two tiny
all-reduces (node totals, split keys).

xgboost semantics (second order, L2 ``reg_lambda``, ``min_child_weight``, learning rate) as used at
scripts/model_comp.py:7 and cml_targetaml_seanalysis.Rmd:1164-1186; trees are complete binary trees of depth
``max_depth`` stored level by level.  CPU tensors (and ``backend="unfused"``) run the same algorithm with torch ops and
NCCL / gloo collectives - the oracle of the tests."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from .. import _native
from ..parallel import fused as F
from ..parallel.dist import all_reduce_


def quantile_bin(x_local: torch.Tensor, nbins: int = 256, sample: int = 200000):
    """Quantile cut points from (a subsample of) this rank's rows, averaged across ranks, and the uint8 bins
    [p, n_pad] of the local rows.  Returns (bins, cuts [p, nbins-1]).

    On the kernel path the cut points are a ``models.trees.Binner``: rank 0 sorts (a strided sample of) each gene's values
    in shared memory (``cml_quantile_edges`` - torch.quantile cannot take a 125k x 20k shard), its edge table is broadcast,
    and every rank bins its own rows against it (``cml_apply_bins``), so all ranks share one binning."""
    n, p = x_local.shape
    if _native.on_device(x_local) and n > 256:
        from .trees import Binner
        world = dist.get_world_size() if dist.is_initialized() else 1
        if world == 1 or dist.get_rank() == 0:
            binner, bins = Binner.fit(x_local, max_bin=nbins)
            edges, n_edges = binner.edges, binner.n_edges
        else:
            edges = torch.empty(p, 256, dtype=torch.float32, device=x_local.device)
            n_edges = torch.empty(p, dtype=torch.int32, device=x_local.device)
        if world > 1:
            dist.broadcast(edges, src=0)
            dist.broadcast(n_edges, src=0)
            binner = Binner(edges, n_edges, max(2, int(n_edges.max().item())))
            bins = binner.transform(x_local)
        return bins, binner
    sub = x_local if n <= sample else x_local[(torch.arange(sample, device=x_local.device) * n) // sample]    # deterministic stride
    qs = torch.linspace(0, 1, nbins + 1, device=x_local.device)[1:-1]
    cuts = torch.quantile(sub.float(), qs, dim=0).t().contiguous()            # [p, nbins-1]
    all_reduce_(cuts)
    cuts /= dist.get_world_size() if dist.is_initialized() else 1
    n_pad = (n + 3) // 4 * 4
    bins = torch.zeros(p, n_pad, dtype=torch.uint8, device=x_local.device)
    bins[:, :n] = torch.searchsorted(cuts, x_local.float().t().contiguous(), right=False).to(torch.uint8)
    return bins, cuts


def apply_cuts(x: torch.Tensor, cuts) -> torch.Tensor:
    if hasattr(cuts, "transform"):                 # a Binner (kernel path)
        return cuts.transform(x)
    n, p = x.shape
    n_pad = (n + 3) // 4 * 4
    bins = torch.zeros(p, n_pad, dtype=torch.uint8, device=x.device)
    bins[:, :n] = torch.searchsorted(cuts, x.float().t().contiguous(), right=False).to(torch.uint8)
    return bins


@dataclass
class LevelTree:
    feat: torch.Tensor      # [2^D - 1] split feature per internal node (level order), -1 = no split (leaf above the last level)
    thr: torch.Tensor       # [2^D - 1] bin threshold (go left iff bin <= thr)
    value: torch.Tensor     # [2^(D+1) - 1] leaf value at every node (used where the walk stops)


@dataclass
class GBMLarge:
    n_estimators: int = 10
    max_depth: int = 3
    learning_rate: float = 0.3
    reg_lambda: float = 1.0
    gamma: float = 0.0
    min_child_weight: float = 1.0
    nbins: int = 256
    backend: str = "auto"                 # "fused" (CUDA kernels, in-kernel reduce-scatter) | "unfused" (torch + all_reduce; CPU path)
    sibling_subtraction: Optional[bool] = None   # build only the smaller child of every split, derive its sibling from the parent
                                          # (halves the samples histogrammed per level).  None = on for the kernel path,
                                          # off for the torch path.
    trees: List[LevelTree] = field(default_factory=list, repr=False)
    cuts: torch.Tensor = field(default=None, repr=False)
    importance_gain_: torch.Tensor = field(default=None, repr=False)
    level_ms_: list = field(default_factory=list, repr=False)

    def fit(self, x_local: Optional[torch.Tensor], y_local: torch.Tensor, bins: Optional[torch.Tensor] = None,
            cuts: Optional[torch.Tensor] = None) -> "GBMLarge":
        if bins is None:
            bins, cuts = quantile_bin(x_local, self.nbins)
        self.cuts = cuts
        backend = self.backend if self.backend != "auto" else ("fused" if _native.on_device(bins) else "unfused")
        if backend == "fused":
            return self._fit_fused(y_local, bins)
        return self._fit_unfused(y_local, bins, backend)

    # ------------------------------------------------------------------------------------------ kernel path
    def _fit_fused(self, y_local: torch.Tensor, bins: torch.Tensor) -> "GBMLarge":
        """Level loop without host synchronisation (see the module docstring).  ``sibling_subtraction`` defaults to on here."""
        dev = y_local.device
        lib = _native.cuda()
        P = _native.ptr
        p, n_pad = bins.shape
        n = y_local.numel()
        D = self.max_depth
        sib = self.sibling_subtraction is not False
        max_nodes = 2 ** max(D - 1, 0)
        max_slots = max(1, max_nodes // 2) if (sib and D > 1) else max_nodes
        H = F.HistEngine(p, self.nbins, max_slots, max_nodes, dev)
        y = y_local.float().contiguous()
        margin = torch.zeros(n, device=dev)
        gh = torch.empty(n, 2, device=dev)
        node = torch.empty(n, dtype=torch.int16, device=dev)
        tot = torch.empty(max(2 ** D, 1), 3, device=dev)
        self.trees = []
        self.importance_gain_ = torch.zeros(p, device=dev)
        self.samples_histogrammed_ = None            # (counting would need a host read-back per level)
        st = _native.stream_ptr
        for _ in range(self.n_estimators):
            _native.check(lib.cml_gbm_grad(P(margin), P(y), P(gh), n, st()), "cml_gbm_grad")
            node.zero_()
            feat_all = torch.full((2 ** D - 1,), -1, dtype=torch.int64, device=dev)
            thr_all = torch.zeros(2 ** D - 1, dtype=torch.int64, device=dev)
            value_all = torch.zeros(2 ** (D + 1) - 1, device=dev)
            for level in range(D + 1):
                n_nodes = 2 ** level
                base = n_nodes - 1
                t = tot[:n_nodes]
                _native.check(lib.cml_gbm_totals(P(node), P(gh), n, n_nodes, P(t), st()), "cml_gbm_totals")
                all_reduce_(t)
                value = (-self.learning_rate * t[:, 0] / (t[:, 1] + self.reg_lambda)).contiguous()
                value_all[base:base + n_nodes] = value
                if level == D:
                    _native.check(lib.cml_gbm_route(P(bins), n_pad, P(node), P(feat_all), P(thr_all), P(value), P(margin), n, 1, st()),
                                  "cml_gbm_route")
                    break
                count = t[:, 2].contiguous()
                if sib and level > 0:
                    # the child with fewer samples (globally) is histogrammed into slot `pair`; its sibling is derived
                    c2 = count.view(-1, 2)
                    right_small = c2[:, 1] < c2[:, 0]
                    pair = torch.arange(n_nodes // 2, device=dev, dtype=torch.int16)
                    none = torch.full_like(pair, -1)
                    slot_of = torch.stack([torch.where(right_small, none, pair), torch.where(right_small, pair, none)], 1).flatten().contiguous()
                    n_slots = n_nodes // 2
                else:
                    slot_of = torch.arange(n_nodes, device=dev, dtype=torch.int16)
                    n_slots = n_nodes
                H.build(bins, n, node, gh, slot_of, n_slots, 1.0, 0.25)
                full = H.derive(level, n_nodes, slot_of, count)
                key = H.best_keys(full, n_nodes, self.reg_lambda, self.gamma, self.min_child_weight)
                all_reduce_(key, dist.ReduceOp.MAX)             # the global winner of every node, identical on all ranks
                gain, feat, thr = F.decode_split_keys(key)
                feat_all[base:base + n_nodes] = feat
                thr_all[base:base + n_nodes] = thr.clamp_min(0)
                split = feat >= 0
                self.importance_gain_.index_add_(0, feat.clamp_min(0), torch.where(split, gain, torch.zeros_like(gain)))
                _native.check(lib.cml_gbm_route(P(bins), n_pad, P(node), P(feat.contiguous()), P(thr.contiguous()), P(value), P(margin), n, 0,
                                                st()), "cml_gbm_route")
            self.trees.append(LevelTree(feat_all, thr_all, value_all))
        return self

    # ------------------------------------------------------------------------------------------ torch path (oracle)
    def _fit_unfused(self, y_local: torch.Tensor, bins: torch.Tensor, backend: str = "unfused") -> "GBMLarge":
        dev = y_local.device
        p = bins.shape[0]
        n = y_local.numel()
        D = self.max_depth
        H = None
        y = y_local.float()
        margin = torch.zeros(n, device=dev)
        ar = torch.arange(n, device=dev)
        self.trees = []
        self.importance_gain_ = torch.zeros(p, device=dev)
        self.samples_histogrammed_ = 0            # local samples that went through a histogram build (work measure)
        for _ in range(self.n_estimators):
            pr = torch.sigmoid(margin)
            gh = torch.stack([pr - y, pr * (1 - pr)], 1).contiguous()
            node = torch.zeros(n, dtype=torch.int16, device=dev)               # index within the current level
            feat_all = torch.full((2 ** D - 1,), -1, dtype=torch.int64, device=dev)
            thr_all = torch.zeros(2 ** D - 1, dtype=torch.int64, device=dev)
            value_all = torch.zeros(2 ** (D + 1) - 1, device=dev)
            alive = torch.ones(n, dtype=torch.bool, device=dev)
            prev_hist = None
            for level in range(D + 1):
                n_nodes = 2 ** level
                base = n_nodes - 1
                # node totals (tiny): local index_add + all-reduce; the third column counts samples (sibling choice)
                tot = torch.zeros(n_nodes, 3, device=dev)
                idx = node[alive].long()
                tot.index_add_(0, idx, torch.cat([gh[alive], torch.ones(int(alive.sum()), 1, device=dev)], 1))
                all_reduce_(tot)
                value_all[base:base + n_nodes] = -self.learning_rate * tot[:, 0] / (tot[:, 1] + self.reg_lambda)
                if level == D:
                    break
                build_node = torch.where(alive, node, torch.full_like(node, -1))
                derive = None
                if self.sibling_subtraction and prev_hist is not None:
                    # the smaller child of every pair is histogrammed, the larger one is parent - smaller
                    cnt = tot[:, 2].view(-1, 2)
                    small_is_right = cnt[:, 1] < cnt[:, 0]
                    small = torch.stack([~small_is_right, small_is_right], 1).flatten()          # [n_nodes] bool
                    build_node = torch.where(alive & small[node.long()], node, torch.full_like(node, -1))
                    derive = ~small
                self.samples_histogrammed_ += int((build_node >= 0).sum())
                hist = F.hist_unfused(bins, n, build_node, gh, n_nodes, p, self.nbins)
                if derive is not None:
                    parent = prev_hist.repeat_interleave(2, dim=0)                               # parent of node j is j // 2
                    sibling = hist.view(-1, 2, p, self.nbins, 2).flip(1).reshape(n_nodes, p, self.nbins, 2)
                    hist = torch.where(derive.view(-1, 1, 1, 1), parent - sibling, hist)
                    hist = torch.where((tot[:, 2] == 0).view(-1, 1, 1, 1), torch.zeros_like(hist), hist)   # children of unsplit nodes
                if self.sibling_subtraction:
                    prev_hist = hist.clone()
                gain, feat, thr = F.best_splits_torch(hist, self.reg_lambda, self.gamma, self.min_child_weight)
                gain, feat, thr = F.agree_splits(gain, feat, thr)       # one decision for all ranks (see its docstring)
                split = feat >= 0
                feat_all[base:base + n_nodes] = torch.where(split, feat, torch.full_like(feat, -1))
                thr_all[base:base + n_nodes] = thr
                self.importance_gain_.index_add_(0, feat[split], gain[split])
                # route the samples: children of node j are 2j, 2j+1 on the next level; samples of unsplit nodes stop
                nl = node.long()
                f_i = feat[nl]
                goes = alive & (f_i >= 0)
                b_i = bins[f_i.clamp_min(0), ar].long()
                right = (b_i > thr[nl]).to(torch.int16)
                stopped = alive & ~goes
                margin = margin + torch.where(stopped, value_all[base + nl], torch.zeros_like(margin))
                alive = goes
                node = torch.where(goes, 2 * node + right, node)
            # samples that reached the last level
            margin = margin + torch.where(alive, value_all[(2 ** D - 1) + node.long()], torch.zeros_like(margin))
            self.trees.append(LevelTree(feat_all, thr_all, value_all))
        return self

    def decision_function(self, bins: torch.Tensor, n: int) -> torch.Tensor:
        dev = bins.device
        if _native.on_device(bins) and self.trees:
            # one kernel: every sample walks all trees (csrc/gbm.cu gbm_predict_kernel)
            feat = torch.stack([t.feat for t in self.trees]).contiguous()
            thr = torch.stack([t.thr for t in self.trees]).contiguous()
            val = torch.stack([t.value for t in self.trees]).float().contiguous()
            out = torch.empty(n, device=dev)
            P = _native.ptr
            rc = _native.cuda().cml_gbm_predict(P(bins), bins.shape[1], n, len(self.trees), self.max_depth, P(feat), P(thr), P(val), P(out),
                                                _native.stream_ptr())
            _native.check(rc, "cml_gbm_predict")
            return out
        ar = torch.arange(n, device=dev)
        D = self.max_depth
        margin = torch.zeros(n, device=dev)
        for t in self.trees:
            node = torch.zeros(n, dtype=torch.int64, device=dev)
            alive = torch.ones(n, dtype=torch.bool, device=dev)
            for level in range(D):
                base = 2 ** level - 1
                f = t.feat[base + node]
                goes = alive & (f >= 0)
                stopped = alive & ~goes
                margin = margin + torch.where(stopped, t.value[base + node], torch.zeros_like(margin))
                b = bins[f.clamp_min(0), ar].long()
                node = torch.where(goes, 2 * node + (b > t.thr[base + node]).long(), node)
                alive = goes
            margin = margin + torch.where(alive, t.value[(2 ** D - 1) + node], torch.zeros_like(margin))
        return margin

    def predict_proba(self, x: torch.Tensor) -> torch.Tensor:
        return torch.sigmoid(self.decision_function(apply_cuts(x, self.cuts), x.shape[0]))
