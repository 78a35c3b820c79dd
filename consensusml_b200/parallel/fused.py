"""Python faces of the fused compute + collective kernels (csrc/fused.cu, csrc/gemm_tcgen05.cu).

Three paths, each with an un-fused comparator built from torch ops + ``torch.distributed`` collectives that the
tests hold the kernels to:

1. ``hist_allreduce``        - per-level (g, h) histograms, reduced across ranks by NVLS ``multimem.red``
2. ``xtr_reduce_scatter``    - lasso gradient ``X^T r`` reduce-scattered by gene with peer ``red.add``
3. ``gram_allgather``        - kernel-matrix rows against every rank's samples; TMA reads the peers' tiles

Nothing comparable exists in the reference (single-process CPU)."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _native
from ..ops import gemm as G
from . import symm
from .dist import equal_shard

_I, _LL, _F, _P, _U, _SZ = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p, ctypes.c_uint, ctypes.c_size_t
_native.register_signature("cml_xtr_rs_scratch_bytes", [_I, _I], _SZ)
_native.register_signature("cml_xtr_reduce_scatter", [_P, _LL, _I, _I, _P, _I, _P, _P, _P, _I, _I, _I, _LL, _U, _P])
_native.register_signature("cml_wait_signals", [_P, _I, _I, _U, _P])
_native.register_signature("cml_xtr_rs_det_scratch_bytes", [_I, _I, _I], _SZ)
_native.register_signature("cml_xtr_reduce_scatter_det", [_P, _LL, _I, _I, _P, _I, _P, _P, _P, _I, _I, _I, _U, _P])
_native.register_signature("cml_xtr_det_reduce", [_P, _P, _I, _I, _I, _P])
_native.register_signature("cml_prox_allgather", [_P, _I, _I, _P, _F, _F, _F, _F, _P, _P, _P, _I, _I, _P, _P, _U, _I, _P, _P])
_native.register_signature("cml_prox_allgather_enet", [_P, _I, _I, _P, _F, _F, _F, _F, _F, _P, _P, _P, _I, _I, _P, _P, _U, _I, _P, _P])
_native.register_signature("cml_xb_residual", [_P, _LL, _I, _I, _P, _F, _P, _P, _P, _P, _P, _P])
_native.register_signature("cml_hist_scratch_bytes", [_I, _I], _SZ)
_native.register_signature("cml_hist_ft", [_I, _I], _I)
_native.register_signature("cml_hist_allreduce", [_P, _LL, _I, _I, _I, _P, _P, _I, _P, _P, _P, _P, _P, _I, _I, _U, _I, _F, _F, _P])
_native.register_signature("cml_split_score_gain", [_P, _I, _I, _I, _F, _F, _F, _P, _P])
_native.register_signature("cml_split_score_gain_at", [_P, _I, _I, _I, _F, _F, _F, _P, _I, _P])
_native.register_signature("cml_hist_build", [_P, _LL, _I, _I, _I, _P, _P, _I, _P, _P, _P, _P, _P, _P, _I, _I, _I, _U, _I, _F, _F, _P])
_native.register_signature("cml_gbm_grad", [_P, _P, _P, _I, _P])
_native.register_signature("cml_gbm_predict", [_P, _LL, _I, _I, _I, _P, _P, _P, _P, _P])
_native.register_signature("cml_gbm_totals", [_P, _P, _I, _I, _P, _P])
_native.register_signature("cml_hist_derive", [_P, _P, _P, _P, _P, _LL, _I, _P])
_native.register_signature("cml_gbm_route", [_P, _LL, _P, _P, _P, _P, _P, _I, _I, _P])

SIG_GRAD, SIG_BETA, SIG_HIST = 0, 8, 16      # signal-pad word ranges (one word per rank each)


def _ptr_array(ptrs: List[int]):
    return (ctypes.c_void_p * len(ptrs))(*ptrs)


def to_bf16_padded(x: torch.Tensor) -> torch.Tensor:
    """fp32 [n, p] -> bf16 [n, pad8(p)] (zero padded), through the conversion kernel."""
    return G.to_bf16_rows(x.float().contiguous())[0]


# ------------------------------------------------------------------------------------------------ path 2
class XtrReduceScatter:
    """``g_owned = sum_ranks (X_r^T R_r)[owned genes]`` with the reduction done by peer atomics inside the kernel.

    ``deterministic=True`` selects the fixed-order form: row chunks are added in chunk order and every source rank stores
    its tile into its own slot at the owner, which adds the slots in rank order (``cml_xtr_det_reduce``) - no float atomics,
    bit-identical from run to run, at the price of a world-times larger landing buffer and one small extra kernel."""

    def __init__(self, p: int, nb: int, device, deterministic: bool = False, n_rows: Optional[int] = None):
        self.p, self.nb, self.deterministic = p, nb, deterministic
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.genes_per_rank = equal_shard(p, self.world)
        lib = _native.cuda()
        if deterministic:
            if n_rows is None:
                raise ValueError("the deterministic reduce-scatter sizes its chunk slices from the local row count: pass n_rows")
            self.n_rows = n_rows
            self.slots = symm.alloc((self.world * self.genes_per_rank * nb,), torch.float32, device)      # [world][genes_per_rank][nb]
            self.grad = symm.alloc((self.genes_per_rank * nb,), torch.float32, device)
            self.scratch = torch.zeros(lib.cml_xtr_rs_det_scratch_bytes(n_rows, p, nb), dtype=torch.uint8, device=device)
        else:
            self.grad = symm.alloc((self.genes_per_rank * nb,), torch.float32, device)       # owner-side accumulator
            self.scratch = torch.zeros(lib.cml_xtr_rs_scratch_bytes(p, nb), dtype=torch.uint8, device=device)
        self.epoch = 0

    @property
    def gene0(self) -> int:
        return self.rank * self.genes_per_rank

    @property
    def genes_owned(self) -> int:
        return max(0, min(self.p, self.gene0 + self.genes_per_rank) - self.gene0)

    def launch(self, x_bf16: torch.Tensor, R: torch.Tensor) -> int:
        """Start the fused kernel for this rank's rows; returns the epoch to wait for."""
        self.epoch += 1
        n, ldx = x_bf16.shape
        if self.deterministic:
            assert n == self.n_rows, "deterministic mode was sized for a different local row count"
            rc = _native.cuda().cml_xtr_reduce_scatter_det(
                _native.ptr(x_bf16), ldx, n, self.p, _native.ptr(R), self.nb, _native.ptr(self.scratch), _ptr_array(self.slots.peer_ptrs),
                _ptr_array(self.slots.signal_ptrs), self.rank, self.world, self.genes_per_rank, self.epoch, _native.stream_ptr())
            _native.check(rc, "cml_xtr_reduce_scatter_det")
            return self.epoch
        rc = _native.cuda().cml_xtr_reduce_scatter(
            _native.ptr(x_bf16), ldx, n, self.p, _native.ptr(R), self.nb, _native.ptr(self.scratch),
            _ptr_array(self.grad.peer_ptrs), _ptr_array(self.grad.signal_ptrs), self.rank, self.world, self.genes_per_rank, 0,
            self.epoch, _native.stream_ptr())
        _native.check(rc, "cml_xtr_reduce_scatter")
        return self.epoch

    def wait(self, epoch: Optional[int] = None) -> None:
        buf = self.slots if self.deterministic else self.grad
        rc = _native.cuda().cml_wait_signals(ctypes.c_void_p(buf.signal_ptrs[self.rank]), SIG_GRAD, self.world,
                                             self.epoch if epoch is None else epoch, _native.stream_ptr())
        _native.check(rc, "cml_wait_signals")
        if self.deterministic:            # add the per-source slots in rank order
            rc = _native.cuda().cml_xtr_det_reduce(_native.ptr(self.slots.local), _native.ptr(self.grad.local), self.genes_per_rank, self.nb,
                                                   self.world, _native.stream_ptr())
            _native.check(rc, "cml_xtr_det_reduce")

    def result(self) -> torch.Tensor:
        """Owned slice [genes_owned, nb] (valid after ``wait``); the caller clears it before the epoch after next."""
        return self.grad.local[: self.genes_owned * self.nb].view(self.genes_owned, self.nb)


def xtr_reduce_scatter_unfused(x: torch.Tensor, R: torch.Tensor, p: int) -> torch.Tensor:
    """Comparator: library matmul + ``reduce_scatter_tensor`` (NCCL / gloo)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    gpr = equal_shard(p, world)
    full = torch.zeros(gpr * world, R.shape[1], dtype=torch.float32, device=x.device)
    full[:p] = x.float().t() @ R.float()
    if world == 1:
        return full[:p]
    out = torch.empty(gpr, R.shape[1], dtype=torch.float32, device=x.device)
    if x.is_cuda:
        dist.reduce_scatter_tensor(out, full)
    else:
        dist.all_reduce(full)
        r = dist.get_rank()
        out = full[r * gpr:(r + 1) * gpr]
    r = dist.get_rank()
    return out[: max(0, min(p, (r + 1) * gpr) - r * gpr)]


# ------------------------------------------------------------------------------------------------ path 3
def gram_allgather(x_local: torch.Tensor, *, kernel: str = "radial", gamma: Optional[float] = None, n_total: Optional[int] = None,
                   shift: float = 0.0, cache: Optional[dict] = None, staged: Optional[bool] = None) -> torch.Tensor:
    """Rows of the kernel matrix for this rank's samples against ALL samples: ``K[local, :]`` of shape
    [n_local, n_total].  Each rank publishes its bf16 block (+ squared norms) in symmetric memory.  With several ranks
    the GEMM kernel itself performs the all-gather: a dozen of its CTAs stream the peers' blocks over NVLink into a local
    staging buffer (bulk async copies) and flag each tile as it lands, while the remaining CTAs run tcgen05 tiles on
    whatever has arrived - own block first, then the peers in ring order.  (Reading peer tiles directly from the TMA
    producer also works and is kept for one rank / ``staged=False``, but peer reads bypass the local L2, so every B tile
    then crosses NVLink once per M-tile.)"""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    n_local, p = x_local.shape
    dev = x_local.device
    ld = (p + 7) // 8 * 8
    n_total = n_local if n_total is None else n_total
    rows = n_local if world == 1 else equal_shard(n_total, world)
    # the symmetric block + norm buffers are allocated (and rendezvoused) once per shape and reused: a rendezvous per call
    # would cost more than the GEMM at small sizes.  Safe to reuse because every call ends with a barrier (below).
    key = (world, rows, ld, str(dev))
    if key not in _GRAM_BUFFERS:
        _GRAM_BUFFERS[key] = (symm.alloc((rows, ld), torch.bfloat16, dev), symm.alloc((rows,), torch.float32, dev))
    xb, nrm = _GRAM_BUFFERS[key]
    if n_local < rows:
        xb.local[n_local:].zero_()
        nrm.local[n_local:].zero_()
    G.to_bf16_rows(x_local.float().contiguous(), shift=shift, out=xb.local[:n_local], norm_out=nrm.local[:n_local])
    _native.synchronize()
    xb.barrier()
    gamma = 1.0 / p if gamma is None else gamma
    epi, alpha = (G.EPI_RBF, gamma) if kernel == "radial" else (G.EPI_SCALE, 1.0)
    use_stage = (world > 1) if staged is None else staged
    if use_stage:
        st = gram_stage_buffers(world, rows, ld, dev)
        if symm.poison_enabled():              # nothing of an earlier call may be read: tiles become valid only through this epoch's flags
            symm.poison_(st["stage"])
            symm.poison_(st["norm"])
        out = G.gemm_tn_staged(xb.local[:n_local], xb.peer_ptrs, n_total, rows, ld, st["stage"], st["norm"], st["flags"],
                               st.next_epoch(), rank, epilogue=epi, alpha=alpha, norm_a=nrm.local[:n_local],
                               peer_norm_ptrs=nrm.peer_ptrs)
        if cache is not None:
            cache["stage"] = st
    else:
        out = G.gemm_tn(xb.local[:n_local], None, n_total, epilogue=epi, alpha=alpha, norm_a=nrm.local[:n_local],
                        b_ptrs=xb.peer_ptrs, norm_b_ptrs=nrm.peer_ptrs, rows_per_src=rows, ldb=ld)
    _native.synchronize()
    xb.barrier()                     # nobody frees / rewrites its block while a peer may still be reading it
    if cache is not None:
        cache["xb"], cache["nrm"] = xb, nrm
    return out


class _StageBuffers(dict):
    def next_epoch(self) -> int:
        self["epoch"] = self.get("epoch", 0) + 1
        return self["epoch"]


_STAGE_CACHE: dict = {}
_GRAM_BUFFERS: dict = {}


def gram_stage_buffers(world: int, rows: int, ld: int, device) -> "_StageBuffers":
    """Local staging buffer, norm buffer and tile flags for the staged all-gather GEMM (cached per shape)."""
    key = (world, rows, ld, str(device))
    if key not in _STAGE_CACHE:
        nflags = _native.cuda().cml_gemm_staged_flags(world, rows)
        _STAGE_CACHE[key] = _StageBuffers(stage=torch.empty(world * rows, ld, dtype=torch.bfloat16, device=device),
                                          norm=torch.empty(world * rows, dtype=torch.float32, device=device),
                                          flags=torch.zeros(nflags, dtype=torch.int32, device=device))
    return _STAGE_CACHE[key]


def gram_allgather_unfused(x_local: torch.Tensor, *, kernel: str = "radial", gamma: Optional[float] = None, shift: float = 0.0):
    """Comparator: NCCL all-gather of the bf16-rounded blocks, then a plain fp32 matmul + epilogue in torch."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    xb, nrm = G.to_bf16_rows(x_local.float().contiguous(), shift=shift)
    if world > 1:
        parts = [torch.empty_like(xb) for _ in range(world)]
        dist.all_gather(parts, xb)
        allx = torch.cat(parts)
    else:
        allx = xb
    a, b = xb.float(), allx.float()
    dot = a @ b.t()
    if kernel != "radial":
        return dot
    gamma = 1.0 / x_local.shape[1] if gamma is None else gamma
    d2 = (a * a).sum(1)[:, None] + (b * b).sum(1)[None, :] - 2 * dot
    return torch.exp(-gamma * d2.clamp_min(0))


# ------------------------------------------------------------------------------------------------ path 1
class HistAllReduce:
    """Per-level gradient histograms ``hist[node, feature, bin, (g, h)]`` summed over ranks inside the build kernel."""

    def __init__(self, p: int, nbins: int, max_nodes: int, device):
        self.p, self.nbins, self.max_nodes = p, nbins, max_nodes
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        size = max_nodes * p * nbins * 2
        self.hist = symm.alloc((size,), torch.float32, device)
        self.local = torch.zeros(size, dtype=torch.float32, device=device)
        self.counters = torch.zeros(p + 16, dtype=torch.int32, device=device)
        self.epoch = 0

    def build(self, bins: torch.Tensor, n: int, node: torch.Tensor, gh: torch.Tensor, n_nodes: int, use_multicast: bool = True,
              g_bound: Optional[float] = None, h_bound: Optional[float] = None):
        """Launch build + reduce; returns the reduced histogram view [n_nodes, p, nbins, 2] after waiting.
        ``g_bound`` / ``h_bound`` bound |g| and |h| (1 and 0.25 for logistic loss); they fix the fixed-point scale of the
        shared-memory accumulation and are measured from ``gh`` when omitted."""
        if g_bound is None or h_bound is None:
            mx = gh.abs().amax(0)
            g_bound, h_bound = float(mx[0]) + 1e-30, float(mx[1]) + 1e-30
        self.epoch += 1
        mc = self.hist.multicast_ptr if (use_multicast and self.world > 1) else 0
        rc = _native.cuda().cml_hist_allreduce(
            _native.ptr(bins), bins.shape[1], n, self.p, self.nbins, _native.ptr(node), _native.ptr(gh), n_nodes,
            _native.ptr(self.local), _native.ptr(self.counters), ctypes.c_void_p(mc), _ptr_array(self.hist.peer_ptrs),
            _ptr_array(self.hist.signal_ptrs), self.rank, self.world, self.epoch, SIG_HIST, float(g_bound), float(h_bound),
            _native.stream_ptr())
        _native.check(rc, "cml_hist_allreduce")
        rc = _native.cuda().cml_wait_signals(ctypes.c_void_p(self.hist.signal_ptrs[self.rank]), SIG_HIST, self.world, self.epoch,
                                             _native.stream_ptr())
        _native.check(rc, "cml_wait_signals")
        return self.hist.local[: n_nodes * self.p * self.nbins * 2].view(n_nodes, self.p, self.nbins, 2)

    def clear(self, n_nodes: int) -> None:
        """Zero the reduced histogram for the next level - after every rank has finished reading it."""
        _native.synchronize()
        self.hist.barrier()
        self.hist.local[: n_nodes * self.p * self.nbins * 2].zero_()
        _native.synchronize()
        self.hist.barrier()


class HistEngine:
    """Per-level gradient histograms with the reduce-scatter-by-feature landing.

    Rank r owns features [r * fpr, (r + 1) * fpr): every rank's build kernel adds its finished feature tiles into the
    OWNER's buffer only (peer ``red.add.v4.f32`` from inside the kernel), so a rank receives 1 / world of the histogram,
    scores only its own features, and the ranks agree on the split with one MAX all-reduce of n_nodes 64-bit keys.
    Compared with the all-reduce landing (``HistAllReduce``) that is 1 / world of the bytes landed per rank and
    1 / world of the split-scoring work.  Slots: a launch builds ``n_slots`` histograms selected by a node -> slot table,
    which is how sibling subtraction (build the smaller child only) and node batching (more open nodes than shared
    memory holds) are expressed.  The landing buffer is cleaned by ``derive`` as it reads, so there is no separate clear
    and no barrier between levels."""

    MAX_SLOTS_PER_LAUNCH = 64

    def __init__(self, p: int, nbins: int, max_slots: int, max_nodes: int, device):
        self.p, self.nbins, self.max_slots = p, nbins, max_slots
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.fpr = ((p + self.world - 1) // self.world + 7) // 8 * 8          # features per rank: a multiple of every tile width
        self.feat0 = self.rank * self.fpr
        self.cells = self.fpr * nbins * 2
        self.landing = symm.alloc((max_slots * self.cells,), torch.float32, device)
        per_launch = min(max_slots, self.MAX_SLOTS_PER_LAUNCH)
        self.local = torch.zeros(per_launch * p * nbins * 2, dtype=torch.float32, device=device)
        self.counters = torch.zeros(p + 16, dtype=torch.int32, device=device)
        self.levels = [torch.zeros(max_nodes * self.cells, dtype=torch.float32, device=device) for _ in range(2)]
        self.best = torch.zeros(max_nodes, dtype=torch.int64, device=device)
        self.epoch = 0
        self.device = device

    def build(self, bins: torch.Tensor, n: int, node: torch.Tensor, gh: torch.Tensor, slot_of: torch.Tensor, n_slots: int,
              g_bound: float, h_bound: float) -> None:
        """Histogram the samples of the nodes that have a slot and land every feature tile on its owner.  Slots beyond what
        shared memory holds per launch are built in batches (each batch re-reads the bins of its own samples only)."""
        lib = _native.cuda()
        S = self.MAX_SLOTS_PER_LAUNCH
        for s0 in range(0, n_slots, S):
            ns = min(S, n_slots - s0)
            so = slot_of if n_slots <= S else torch.where((slot_of >= s0) & (slot_of < s0 + ns), slot_of - s0, torch.full_like(slot_of, -1))
            self.epoch += 1
            off = s0 * self.cells * 4
            rc = lib.cml_hist_build(_native.ptr(bins), bins.shape[1], n, self.p, self.nbins, _native.ptr(node), _native.ptr(gh), ns,
                                    _native.ptr(so.contiguous()), _native.ptr(self.local), _native.ptr(self.counters), None,
                                    _ptr_array([a + off for a in self.landing.peer_ptrs]), _ptr_array(self.landing.signal_ptrs), self.rank,
                                    self.world, self.fpr, self.epoch, SIG_HIST, float(g_bound), float(h_bound), _native.stream_ptr())
            _native.check(rc, "cml_hist_build")
            rc = lib.cml_wait_signals(ctypes.c_void_p(self.landing.signal_ptrs[self.rank]), SIG_HIST, self.world, self.epoch,
                                      _native.stream_ptr())
            _native.check(rc, "cml_wait_signals")

    def derive(self, level: int, n_nodes: int, slot_of: torch.Tensor, count: torch.Tensor) -> torch.Tensor:
        """Full histograms of this level's nodes for the owned features, [n_nodes, fpr, nbins, 2]: built slots are copied,
        derived nodes are parent - sibling (the parent level is the buffer ``derive`` filled one level up)."""
        full, parent = self.levels[level & 1], self.levels[(level + 1) & 1]
        rc = _native.cuda().cml_hist_derive(_native.ptr(self.landing.local), _native.ptr(parent) if n_nodes > 1 else None, _native.ptr(full),
                                            _native.ptr(slot_of.contiguous()), _native.ptr(count.contiguous()), self.cells, n_nodes,
                                            _native.stream_ptr())
        _native.check(rc, "cml_hist_derive")
        return full[: n_nodes * self.cells].view(n_nodes, self.fpr, self.nbins, 2)

    def best_keys(self, full: torch.Tensor, n_nodes: int, lam: float, gamma: float, min_child_weight: float) -> torch.Tensor:
        """Packed (gain | feature | bin) key of the best split of every node over the OWNED features (global feature ids)."""
        rc = _native.cuda().cml_split_score_gain_at(_native.ptr(full), n_nodes, self.fpr, self.nbins, lam, gamma, min_child_weight,
                                                    _native.ptr(self.best), self.feat0, _native.stream_ptr())
        _native.check(rc, "cml_split_score_gain_at")
        return self.best[:n_nodes]


def decode_split_keys(key: torch.Tensor):
    """(gain fp32, feature int64, bin int64) from the packed keys; -1 / 0 where a node has no valid split."""
    has = key != 0
    gain = ((key >> 32) & 0xFFFFFFFF).to(torch.int32).view(torch.float32)
    feat = 0xFFFFF - ((key >> 12) & 0xFFFFF)
    b = 0xFFF - (key & 0xFFF)
    neg = -torch.ones_like(feat)
    return torch.where(has, gain, torch.zeros_like(gain)), torch.where(has, feat, neg), torch.where(has, b, neg)


def hist_unfused(bins: torch.Tensor, n: int, node: torch.Tensor, gh: torch.Tensor, n_nodes: int, p: int, nbins: int) -> torch.Tensor:
    """Comparator: histogram by ``index_add_`` + ``all_reduce``."""
    dev = bins.device
    hist = torch.zeros(n_nodes * p * nbins, 2, dtype=torch.float32, device=dev)
    valid = node[:n] >= 0
    idx_s = torch.nonzero(valid).flatten()
    nd = node[:n][idx_s].long()
    for f in range(p):
        flat = (nd * p + f) * nbins + bins[f, :n][idx_s].long()
        hist.index_add_(0, flat, gh[idx_s])
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(hist)
    return hist.view(n_nodes, p, nbins, 2)


def best_splits(hist: torch.Tensor, lam: float = 1.0, gamma: float = 0.0, min_child_weight: float = 1.0):
    """Per open node: (gain, feature, bin) from the reduced histogram via ``cml_split_score_gain``.
    Returns int64 tensors feature, bin (-1 where no valid split) and fp32 gain."""
    n_nodes, p, nbins, _ = hist.shape
    best = torch.zeros(n_nodes, dtype=torch.int64, device=hist.device)
    rc = _native.cuda().cml_split_score_gain(_native.ptr(hist.contiguous()), n_nodes, p, nbins, lam, gamma, min_child_weight,
                                             _native.ptr(best), _native.stream_ptr())
    _native.check(rc, "cml_split_score_gain")
    key = best
    has = key != 0
    gain = ((key >> 32) & 0xFFFFFFFF).to(torch.int32).view(torch.float32)
    feat = 0xFFFFF - ((key >> 12) & 0xFFFFF)
    b = 0xFFF - (key & 0xFFF)
    return torch.where(has, gain, torch.zeros_like(gain)), torch.where(has, feat, -torch.ones_like(feat)), \
        torch.where(has, b, -torch.ones_like(b))


def agree_splits(gain: torch.Tensor, feat: torch.Tensor, thr: torch.Tensor):
    """Make the split decision single-sourced across ranks.

    Every rank scores its own copy of the reduced histogram.  The copies are sums of floats that arrive in a different
    order on every rank, so they may differ in the last ulp and a near-tie in gain could otherwise pick a different
    (feature, bin) on different ranks - which would route samples inconsistently without any error.  The decision is
    therefore packed into one ordered 64-bit key per node (gain bits | inverted feature | inverted bin, the same key
    ``cml_split_score_gain`` builds) and MAX-all-reduced: n_nodes words per level."""
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return gain, feat, thr
    has = feat >= 0
    gbits = gain.float().contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF      # gains are > 0: bit order = value order
    key = (gbits << 32) | ((0xFFFFF - feat.clamp_min(0)) << 12) | (0xFFF - thr.clamp_min(0))
    key = torch.where(has, key, torch.zeros_like(key))
    dist.all_reduce(key, op=dist.ReduceOp.MAX)
    has = key != 0
    g = ((key >> 32) & 0xFFFFFFFF).to(torch.int32).view(torch.float32)
    f = 0xFFFFF - ((key >> 12) & 0xFFFFF)
    b = 0xFFF - (key & 0xFFF)
    neg = -torch.ones_like(f)
    return torch.where(has, g, torch.zeros_like(g)), torch.where(has, f, neg), torch.where(has, b, neg)


def best_splits_torch(hist: torch.Tensor, lam: float = 1.0, gamma: float = 0.0, min_child_weight: float = 1.0):
    """Same contract as ``best_splits`` in plain torch (CPU path and oracle): xgboost's second-order gain
    ``G_L^2/(H_L+lam) + G_R^2/(H_R+lam) - G^2/(H+lam) - gamma`` over every cut after an occupied bin (ties: lowest feature, bin)."""
    hh = hist.double()
    cg, ch = hh[..., 0].cumsum(-1), hh[..., 1].cumsum(-1)
    tg, th = cg[..., -1:], ch[..., -1:]
    gain = cg ** 2 / (ch + lam) + (tg - cg) ** 2 / (th - ch + lam) - tg ** 2 / (th + lam) - gamma     # as the kernel: not halved
    valid = (hh[..., 1] > 0) & (ch >= min_child_weight) & (th - ch >= min_child_weight) & (th - ch > 0)
    valid[..., -1] = False
    gain = torch.where(valid, gain, torch.full_like(gain, -1.0))
    n_nodes, p, nbins = gain.shape
    flat = gain.reshape(n_nodes, -1)
    best, arg = flat.max(1)
    has = best > 0
    feat, b = arg // nbins, arg % nbins
    neg = -torch.ones_like(feat)
    return torch.where(has, best, torch.zeros_like(best)).float(), torch.where(has, feat, neg), torch.where(has, b, neg)


# ------------------------------------------------------------------------------------------------ path 2b
# Replicate-batched gradient as a tcgen05 GEMM fused with its reduce-scatter (csrc/gemm_rs.cu).  Executed on the host model
# of the asynchronous units (tests/test_emu_gemm.py) and by models/lasso_large.BatchedLassoLarge; its first hardware run is
# tests/test_yy_gemm_rs_gpu.py.
SIG_GEMM_RS = 32
_native.register_signature("cml_xtr_gemm_reduce_scatter",
                           [ctypes.c_int] * 4 + [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p,
                                                 ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                                 ctypes.c_uint, ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_void_p])


def split_bf16(r: torch.Tensor) -> torch.Tensor:
    """[B, K] fp32 -> [2B, K] bf16: rows [0, B) the bf16 rounding of r, rows [B, 2B) the rounding of the remainder, so
    that hi + lo carries 16 mantissa bits into a bf16 tensor-core product."""
    hi = r.bfloat16()
    lo = (r - hi.float()).bfloat16()
    return torch.cat([hi, lo], 0).contiguous()


class XtrGemmReduceScatter:
    """``G_owned[genes_owned, B] = sum_ranks (Xt_r . Rt_r^T)[owned genes]`` - GEMM and reduce-scatter in one kernel.

    ``xt`` is the gene-major bf16 copy of this rank's rows ([p, ld], ld % 8 == 0), ``rt`` the replicate-major bf16
    residuals ([B, ld] or the [2B, ld] hi/lo split).  Results alternate between two symmetric buffers; the one returned by
    ``result()`` stays valid until the next ``launch``."""

    def __init__(self, p: int, B: int, device, use_tma_reduce: bool = False, *, rank: Optional[int] = None, world: Optional[int] = None,
                 bufs=None):
        """``rank`` / ``world`` / ``bufs`` (two SymmBuffers of shape ``(genes_per_rank(p, world), B)``) override the process
        group: that is how several ranks are played in one process (``symm.LocalGroup``)."""
        assert B % 32 == 0, "B must be a multiple of 32"
        self.world = world if world is not None else (dist.get_world_size() if dist.is_initialized() else 1)
        self.rank = rank if rank is not None else (dist.get_rank() if dist.is_initialized() else 0)
        self.p, self.B, self.use_tma = p, B, use_tma_reduce
        self.genes_per_rank = self.genes_per_rank_for(p, self.world)
        self.gene0 = min(p, self.rank * self.genes_per_rank)
        self.genes_owned = max(0, min(p, self.gene0 + self.genes_per_rank) - self.gene0)
        self.bufs = bufs if bufs is not None else [symm.alloc((self.genes_per_rank, B), torch.float32, device) for _ in range(2)]
        self.done = torch.zeros(4, dtype=torch.int32, device=device)
        self.epoch = 0

    @staticmethod
    def genes_per_rank_for(p: int, world: int) -> int:
        return ((p + world - 1) // world + 31) // 32 * 32

    def launch(self, xt: torch.Tensor, rt: torch.Tensor, alpha: float = 1.0) -> None:
        assert xt.dtype == torch.bfloat16 and rt.dtype == torch.bfloat16 and xt.shape[1] == rt.shape[1] and xt.shape[1] % 8 == 0
        assert xt.shape[0] == self.p and rt.shape[0] in (self.B, 2 * self.B)
        self.epoch += 1
        buf, nxt = self.bufs[self.epoch & 1], self.bufs[(self.epoch + 1) & 1]
        nxt.local.zero_()            # the buffer of the NEXT epoch: peers cannot add to it before they see this epoch's signal
        rc = _native.cuda().cml_xtr_gemm_reduce_scatter(
            self.p, rt.shape[0], self.B, xt.shape[1], _native.ptr(xt), xt.stride(0), _native.ptr(rt), rt.stride(0),
            _ptr_array(buf.peer_ptrs), self.B, self.genes_per_rank, _ptr_array(buf.signal_ptrs), self.rank, self.world,
            _native.ptr(self.done), self.epoch, SIG_GEMM_RS, float(alpha), int(self.use_tma), _native.stream_ptr())
        _native.check(rc, "cml_xtr_gemm_reduce_scatter")

    def wait(self) -> None:
        buf = self.bufs[self.epoch & 1]
        rc = _native.cuda().cml_wait_signals(ctypes.c_void_p(buf.signal_ptrs[self.rank]), SIG_GEMM_RS, self.world, self.epoch,
                                             _native.stream_ptr())
        _native.check(rc, "cml_wait_signals")

    def result(self) -> torch.Tensor:
        return self.bufs[self.epoch & 1].local[: self.genes_owned]


def xtr_gemm_rs_unfused(xt: torch.Tensor, rt: torch.Tensor, p: int, B: int, genes_per_rank: int) -> torch.Tensor:
    """Comparator / CPU path: fp32 matmul of the same bf16 operands, all-reduce, slice of the owned genes."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    g = xt.float() @ rt.float().t()
    if rt.shape[0] == 2 * B:
        g = g[:, :B] + g[:, B:]
    if world > 1:
        dist.all_reduce(g)
    g0 = min(p, rank * genes_per_rank)
    return g[g0: min(p, g0 + genes_per_rank)]
