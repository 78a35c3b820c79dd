"""Sample-sharded gradient-boosted trees for large cohorts (BASELINE.json config 4: 1M samples x 20k genes).

The small-cohort engine (models/trees.py) grows whole trees inside CTAs; at this size the per-level gradient
histogram is the job: every rank streams its uint8-binned sample shard once per level and the (g, h) histograms of
all open nodes are summed across ranks INSIDE the build kernel (``parallel.fused.HistAllReduce``: shared-memory
int32 accumulation -> rank-local tile -> NVLS ``multimem.red`` into every rank's copy).  Each rank then scores the
splits of the reduced histogram (``cml_split_score_gain``) and - because the histogram is identical everywhere -
reaches the same decision without further communication.

xgboost semantics (second order, L2 ``reg_lambda``, ``min_child_weight``, learning rate) as used at
scripts/model_comp.py:7 and cml_targetaml_seanalysis.Rmd:1164-1186; trees are complete binary trees of depth
``max_depth`` stored level by level."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..parallel import fused as F
from ..parallel.dist import all_reduce_


def quantile_bin(x_local: torch.Tensor, nbins: int = 256, sample: int = 200000):
    """Quantile cut points from (a subsample of) this rank's rows, averaged across ranks, and the uint8 bins
    [p, n_pad] of the local rows.  Returns (bins, cuts [p, nbins-1])."""
    n, p = x_local.shape
    sub = x_local if n <= sample else x_local[torch.randperm(n, device=x_local.device)[:sample]]
    qs = torch.linspace(0, 1, nbins + 1, device=x_local.device)[1:-1]
    cuts = torch.quantile(sub.float(), qs, dim=0).t().contiguous()            # [p, nbins-1]
    all_reduce_(cuts)
    cuts /= dist.get_world_size() if dist.is_initialized() else 1
    n_pad = (n + 3) // 4 * 4
    bins = torch.zeros(p, n_pad, dtype=torch.uint8, device=x_local.device)
    bins[:, :n] = torch.searchsorted(cuts, x_local.float().t().contiguous(), right=False).to(torch.uint8)
    return bins, cuts


def apply_cuts(x: torch.Tensor, cuts: torch.Tensor) -> torch.Tensor:
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
    use_multicast: bool = True
    backend: str = "auto"                 # "fused" (CUDA kernels, in-kernel all-reduce) | "unfused" (torch + all_reduce; CPU path)
    sibling_subtraction: bool = False     # build only the smaller child of every split, derive its sibling from the parent
                                          # (halves the samples histogrammed per level; exact with the kernel's fixed-point sums).
                                          # Opt-in until it has been run on the GPU path.
    trees: List[LevelTree] = field(default_factory=list, repr=False)
    cuts: torch.Tensor = field(default=None, repr=False)
    importance_gain_: torch.Tensor = field(default=None, repr=False)
    level_ms_: list = field(default_factory=list, repr=False)

    def fit(self, x_local: Optional[torch.Tensor], y_local: torch.Tensor, bins: Optional[torch.Tensor] = None,
            cuts: Optional[torch.Tensor] = None) -> "GBMLarge":
        dev = y_local.device
        if bins is None:
            bins, cuts = quantile_bin(x_local, self.nbins)
        self.cuts = cuts
        p = bins.shape[0]
        n = y_local.numel()
        D = self.max_depth
        backend = self.backend if self.backend != "auto" else ("fused" if bins.is_cuda else "unfused")
        H = F.HistAllReduce(p, self.nbins, 2 ** (D - 1) if D > 0 else 1, dev) if backend == "fused" else None
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
                e0 = e1 = None
                if bins.is_cuda:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
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
                if backend == "fused":
                    hist = H.build(bins, n, build_node, gh, n_nodes, use_multicast=self.use_multicast, g_bound=1.0, h_bound=0.25)
                else:
                    hist = F.hist_unfused(bins, n, build_node, gh, n_nodes, p, self.nbins)
                if derive is not None:
                    hist = hist.clone() if backend == "fused" else hist
                    parent = prev_hist.repeat_interleave(2, dim=0)                               # parent of node j is j // 2
                    sibling = hist.view(-1, 2, p, self.nbins, 2).flip(1).reshape(n_nodes, p, self.nbins, 2)
                    hist = torch.where(derive.view(-1, 1, 1, 1), parent - sibling, hist)
                    hist = torch.where((tot[:, 2] == 0).view(-1, 1, 1, 1), torch.zeros_like(hist), hist)   # children of unsplit nodes
                if self.sibling_subtraction:
                    prev_hist = hist.clone()
                if backend == "fused":
                    gain, feat, thr = F.best_splits(hist, self.reg_lambda, self.gamma, self.min_child_weight)
                else:
                    gain, feat, thr = F.best_splits_torch(hist, self.reg_lambda, self.gamma, self.min_child_weight)
                gain, feat, thr = F.agree_splits(gain, feat, thr)       # one decision for all ranks (see its docstring)
                if e1 is not None:
                    e1.record()
                    e1.synchronize()
                    self.level_ms_.append((level, e0.elapsed_time(e1)))
                if backend == "fused":
                    H.clear(n_nodes)
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
