"""C-SVC for cohorts whose kernel matrix cannot exist: working-set decomposition over sample-sharded ranks (K7, large n).

The reference fits e1071::svm on 93 samples (cml_targetaml_seanalysis.Rmd:126-210); the single-CTA SMO of models/svm.py
covers that regime (n <= 10,240: the whole kernel matrix is materialised).  At 10^5 - 10^6 samples K (n x n) is out of
reach, so this solver never forms it:

  every outer iteration
    1. each rank nominates q / world variables of ITS OWN shard: the most violating half from I_up, the other half from
       I_low (the globally most violating pair is always among the nominees, which is what convergence needs);
    2. the nominees' feature rows are published in a symmetric buffer and every rank computes the kernel COLUMNS
       K[local samples, working set] with the all-gather -> tcgen05 Gram kernel (``parallel.fused``: copier CTAs pull the
       peers' rows over NVLink while the GEMM CTAs already run on what has landed; RBF exponent in the epilogue) - this is
       the consumer of fused path 3.  With one rank an LRU cache keeps the columns of recently used variables;
    3. the q x q sub-problem (the gathered rows of those columns) is solved by the warm-started single-CTA SMO kernel
       (``cml_svm_smo_warm``), replicated on every rank - it is deterministic, so all ranks get the same step;
    4. each rank updates its slice of the gradient: G_local += y_local * (K[local, ws] @ (d_alpha * y_ws)).
  until max violation (m - M, all-reduced) < eps.

Collectives per iteration: two scalars (MAX), q floats x 4 (all-gather of y / alpha / G / C of the nominees), q x q floats
(all-gather of the sub-problem matrix) - the sample-feature traffic is inside the fused GEMM.  CPU tensors run the same
algorithm with torch ops (and gloo), which is how the decomposition logic is tested without a GPU."""
from __future__ import annotations

import ctypes
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from .. import _native
from ..ops import gemm as G
from . import svm as S


def _world():
    return (dist.get_world_size(), dist.get_rank()) if dist.is_initialized() else (1, 0)


def _all_gather_cat(t: torch.Tensor) -> torch.Tensor:
    world, _ = _world()
    if world == 1:
        return t
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t.contiguous())
    return torch.cat(parts)


class ColumnCache:
    """LRU cache of kernel columns K[:, j] (single-rank mode): a working set keeps most of its members from one outer
    iteration to the next, so only the columns of new members go through the GEMM."""

    def __init__(self, n_rows: int, capacity: int, device):
        self.buf = torch.empty(capacity, n_rows, dtype=torch.float32, device=device)
        self.slot: "OrderedDict[int, int]" = OrderedDict()
        self.free = list(range(capacity))
        self.hits = self.misses = 0

    def lookup(self, ids):
        """Slots of the cached ids (refreshing their age) and the list of ids that are missing."""
        have, miss = {}, []
        for j in ids:
            if j in self.slot:
                self.slot.move_to_end(j)
                have[j] = self.slot[j]
                self.hits += 1
            else:
                miss.append(j)
                self.misses += 1
        return have, miss

    def insert(self, ids, cols: torch.Tensor, keep) -> dict:
        out = {}
        for k, j in enumerate(ids):
            if not self.free:
                # evict the least recently used entry that the current working set does not need
                for old in self.slot:
                    if old not in keep:
                        self.free.append(self.slot.pop(old))
                        break
            s = self.free.pop()
            self.buf[s] = cols[k]
            self.slot[j] = s
            out[j] = s
        return out


@dataclass
class LargeSVC:
    kernel: str = "radial"
    cost: float = 1.0
    gamma: Optional[float] = None
    tol: float = 1e-3
    q: int = 512                       # working-set size (split evenly over the ranks)
    max_outer: int = 2000
    inner_iter: int = 100000
    cache_columns: int = 4096          # LRU capacity (single rank)

    alpha_: torch.Tensor = field(default=None, repr=False)      # local slice
    rho_: float = 0.0
    outer_iters_: int = 0
    objective_: float = float("nan")
    gap_: float = float("nan")
    cache_hit_rate_: float = float("nan")

    # ---- kernel columns -------------------------------------------------------------------------------------------
    def _columns(self, x_local: torch.Tensor, x_ws_local: torch.Tensor, gamma: float, q_total: int) -> torch.Tensor:
        """K[local samples, all nominees] -> [n_local, q_total]: the nominees' rows of every rank against this rank's samples."""
        world, _ = _world()
        if _native.on_device(x_local):
            from ..parallel import fused as F
            # gram_allgather computes rows = its first argument against the all-gathered blocks: here the "block" a rank
            # publishes is its nominee rows, and the left operand is the local shard
            return self._gram_local_vs_gathered(x_local, x_ws_local, gamma, q_total)
        xw = _all_gather_cat(x_ws_local)
        if self.kernel == "radial":
            return torch.exp(-gamma * torch.cdist(x_local.double(), xw.double()) ** 2).float()
        return (x_local.double() @ xw.double().t()).float()

    def _gram_local_vs_gathered(self, x_local, x_ws_local, gamma, q_total):
        from ..parallel import fused as F
        from ..parallel import symm
        world, rank = _world()
        dev = x_local.device
        qw, p = x_ws_local.shape
        ld = (p + 7) // 8 * 8
        st = self.__dict__.setdefault("_bufs", {})
        if "xb" not in st:
            st["xa"], st["na"] = G.to_bf16_rows(x_local.float().contiguous())          # the local shard in bf16, once
            st["xb"] = symm.alloc((qw, ld), torch.bfloat16, dev)                        # nominee rows, published per iteration
            st["nb"] = symm.alloc((qw,), torch.float32, dev)
            st["stage"] = F.gram_stage_buffers(world, qw, ld, dev) if world > 1 else None
        xb, nb = st["xb"], st["nb"]
        G.to_bf16_rows(x_ws_local.float().contiguous(), out=xb.local, norm_out=nb.local)
        _native.synchronize()
        xb.barrier()
        epi, alpha = (G.EPI_RBF, gamma) if self.kernel == "radial" else (G.EPI_SCALE, 1.0)
        if world > 1:
            sg = st["stage"]
            if symm.poison_enabled():          # debug mode: stale nominee rows of the previous outer iteration must never be read
                symm.poison_(sg["stage"])
                symm.poison_(sg["norm"])
            out = G.gemm_tn_staged(st["xa"], xb.peer_ptrs, q_total, qw, ld, sg["stage"], sg["norm"], sg["flags"], sg.next_epoch(), rank,
                                   epilogue=epi, alpha=alpha, norm_a=st["na"], peer_norm_ptrs=nb.peer_ptrs)
        else:
            out = G.gemm_tn(st["xa"], [xb.local], q_total, epilogue=epi, alpha=alpha, norm_a=st["na"], norm_b=[nb.local])
        _native.synchronize()
        xb.barrier()                  # nobody overwrites its nominee block while a peer may still be reading it
        return out

    # ---- sub-problem ----------------------------------------------------------------------------------------------
    def _solve_sub(self, Kww, y_ws, c_ws, a_ws, g_ws):
        q = y_ws.numel()
        if _native.on_device(Kww):
            Kww = Kww.float().contiguous()
            alpha = torch.empty(1, q, dtype=torch.float32, device=Kww.device)
            rho = torch.empty(1, dtype=torch.float32, device=Kww.device)
            iters = torch.empty(1, dtype=torch.int32, device=Kww.device)
            obj = torch.empty(1, dtype=torch.float32, device=Kww.device)
            P = _native.ptr
            rc = _native.cuda().cml_svm_smo_warm(1, q, P(Kww), 0, Kww.stride(0), P(y_ws.float().contiguous()), 0, P(c_ws.float().contiguous()), q,
                                                 float(self.tol * 0.5), int(self.inner_iter), P(a_ws.float().contiguous()),
                                                 P(g_ws.float().contiguous()), P(alpha), None, P(rho), P(iters), P(obj), _native.stream_ptr())
            _native.check(rc, "cml_svm_smo_warm")
            return alpha[0]
        a, _, _ = S.smo_numpy(Kww.double().numpy(), y_ws.double().numpy(), c_ws.double().numpy(), self.tol * 0.5, self.inner_iter,
                              alpha0=a_ws.double().numpy(), g0=g_ws.double().numpy())
        return torch.from_numpy(a).to(y_ws.dtype)

    # ---- fit ------------------------------------------------------------------------------------------------------
    def fit(self, x_local: torch.Tensor, y_local: torch.Tensor) -> "LargeSVC":
        """x_local: this rank's (already scaled) samples [n_local, p]; y_local in {0, 1}.  Collective over the ranks."""
        world, rank = _world()
        dev = x_local.device
        n_loc, p = x_local.shape
        gamma = self.gamma if self.gamma is not None else 1.0 / p
        y = (2 * y_local.float() - 1).to(dev)
        C = float(self.cost)
        alpha = torch.zeros(n_loc, device=dev)
        Gr = -torch.ones(n_loc, device=dev)
        qw = max(2, (self.q // world) // 2 * 2)
        qw = min(qw, n_loc // 2 * 2)
        q_total = qw * world
        cache = ColumnCache(n_loc, max(self.cache_columns, 2 * qw), dev) if (world == 1 and self.cache_columns > 0) else None
        ninf = torch.full((n_loc,), -float("inf"), device=dev)
        it = 0
        for it in range(1, self.max_outer + 1):
            up = ((y > 0) & (alpha < C)) | ((y < 0) & (alpha > 0))
            low = ((y > 0) & (alpha > 0)) | ((y < 0) & (alpha < C))
            up_s = torch.where(up, -y * Gr, ninf)
            low_s = torch.where(low, y * Gr, ninf)
            mm = torch.stack([up_s.max(), low_s.max()])
            if world > 1:
                dist.all_reduce(mm, op=dist.ReduceOp.MAX)
            self.gap_ = float(mm[0] + mm[1])
            if self.gap_ < self.tol:
                it -= 1
                break
            iu = torch.topk(up_s, qw // 2).indices
            low_s2 = low_s.clone()
            low_s2[iu] = -float("inf")                                  # a free variable can be in both sets: nominate it once
            il = torch.topk(low_s2, qw // 2).indices
            ws = torch.cat([iu, il])
            valid = torch.cat([up_s[iu] > -float("inf"), low_s2[il] > -float("inf")])      # fewer candidates than slots: dead entries
            # ---- kernel columns of the nominees against the local samples
            if cache is not None:
                ids = ws.tolist()
                have, miss = cache.lookup(ids)
                if miss:
                    mi = torch.tensor(miss, device=dev)
                    cols = self._columns_single(x_local, x_local[mi], gamma)              # [n_loc, len(miss)]
                    have.update(cache.insert(miss, cols.t(), set(ids)))
                K_lw = cache.buf[torch.tensor([have[j] for j in ids], device=dev)].t()    # [n_loc, q]
            else:
                K_lw = self._columns(x_local, x_local[ws], gamma, q_total)
            # ---- gather the sub-problem (tiny): labels, bounds, alpha, gradient of the nominees and their q x q kernel block
            c_ws = _all_gather_cat(torch.where(valid, torch.full((qw,), C, device=dev), torch.zeros(qw, device=dev)))
            y_ws = _all_gather_cat(y[ws])
            a_ws = _all_gather_cat(alpha[ws])
            g_ws = _all_gather_cat(Gr[ws])
            Kww = _all_gather_cat(K_lw[ws].contiguous())                                  # rows of rank r's nominees: [q_total, q_total]
            a_new = self._solve_sub(Kww, y_ws, c_ws, a_ws, g_ws).to(alpha.dtype)
            delta = (a_new - a_ws) * y_ws                                                 # y_j d_alpha_j
            Gr += y * (K_lw.to(Gr.dtype) @ delta.to(Gr.dtype))
            mine = a_new[rank * qw:(rank + 1) * qw]
            alpha[ws] = torch.where(valid, mine, alpha[ws])
        self.outer_iters_ = it
        self.alpha_ = alpha
        # rho as libsvm: mean of y G over the free variables, else the midpoint of the bounds
        yg = y * Gr
        free = (alpha > 0) & (alpha < C)
        stats = torch.stack([yg[free].sum(), free.sum().to(yg.dtype)])
        ub = torch.where(((alpha >= C) & (y < 0)) | ((alpha <= 0) & (y > 0)), yg, torch.full_like(yg, float("inf"))).min()
        lb = torch.where(((alpha >= C) & (y > 0)) | ((alpha <= 0) & (y < 0)), yg, torch.full_like(yg, -float("inf"))).max()
        obj = 0.5 * (alpha * (Gr - 1)).sum().reshape(1)
        if world > 1:
            dist.all_reduce(stats)
            dist.all_reduce(obj)
            b = torch.stack([-ub, lb])
            dist.all_reduce(b, op=dist.ReduceOp.MAX)
            ub, lb = -b[0], b[1]
        self.rho_ = float(stats[0] / stats[1]) if float(stats[1]) > 0 else float(0.5 * (ub + lb))
        self.objective_ = float(obj)
        if cache is not None and cache.hits + cache.misses:
            self.cache_hit_rate_ = cache.hits / (cache.hits + cache.misses)
        self._x, self._y = x_local, y
        self._sharded = world > 1               # decision_function is a collective only for a model fitted over several ranks
        self._gamma = gamma
        return self

    def _columns_single(self, x_local, x_sel, gamma):
        if _native.on_device(x_local):
            return (G.rbf_kernel(x_local, x_sel, gamma) if self.kernel == "radial" else G.linear_kernel(x_local, x_sel)).float()
        if self.kernel == "radial":
            return torch.exp(-gamma * torch.cdist(x_local.double(), x_sel.double()) ** 2).float()
        return (x_local.double() @ x_sel.double().t()).float()

    def decision_function(self, x_new: torch.Tensor) -> torch.Tensor:
        """Decision values of new samples: every rank contributes the support vectors of its shard (all-reduced)."""
        coef = self.alpha_ * self._y
        sv = torch.nonzero(coef != 0).flatten()
        if sv.numel():
            Kt = self._columns_single(x_new, self._x[sv], self._gamma)
            dec = Kt.to(coef.dtype) @ coef[sv]
        else:
            dec = torch.zeros(x_new.shape[0], device=x_new.device)
        if self._sharded:
            dist.all_reduce(dec)
        return dec - self.rho_
