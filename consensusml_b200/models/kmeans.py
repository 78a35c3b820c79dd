"""Batched k-means with restarts and the k sweep of BuieRProj/KM_WHATEVER.Rmd:151-161
(``kmeans(df, centers = k, nstart = 25)`` for k = 20..2 with ``set.seed(1000)``), plus the logistic regression of the
risk label on the cluster id (:169-184).

On CUDA every (k, restart) pair of the sweep is one CTA of ONE launch (``kmeans_lloyd_kernel``, csrc/kmeans.cu: k-means++
seeding and Lloyd iterations with the centres in shared memory, no floating-point atomics); cohorts too large for that
shape iterate with the tcgen05 squared-distance GEMM + ``cml_kmeans_assign`` + ``cml_kmeans_update``.  CPU tensors run the
same algorithm in torch.  R's default is Hartigan-Wong; Lloyd iterations from k-means++ seeds with 25 restarts reach the
same minima on data of this size - acceptance is the total within-cluster sum of squares, not the label permutation."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Sequence

import ctypes

import numpy as np
import torch

from .. import _native

_I, _LL, _P, _ULL = ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_ulonglong
_native.register_signature("cml_kmeans_lloyd", [_P, _LL, _I, _I, _P, _I, _I, _I, _I, _ULL, _P, _P, _P, _P, _P, _P, _P])
_native.register_signature("cml_kmeans_update", [_P, _LL, _I, _I, _I, _P, _P, _P, _P])


@dataclass
class KMeansResult:
    centers: torch.Tensor        # [k, d]
    labels: torch.Tensor         # [n]
    tot_withinss: float
    withinss: torch.Tensor       # [k]
    iters: int


def _sqdist(x: torch.Tensor, c: torch.Tensor) -> torch.Tensor:
    """[R, n, k] squared distances for R restarts."""
    xx = (x * x).sum(1)[None, :, None]
    cc = (c * c).sum(2)[:, None, :]
    return (xx + cc - 2 * torch.einsum("nd,rkd->rnk", x, c)).clamp_min(0)


def _kmeans_device(x: torch.Tensor, ks: Sequence[int], nstart: int, max_iter: int, seed: int) -> Dict[int, KMeansResult]:
    """All (k, restart) problems of a sweep in one launch of ``kmeans_lloyd_kernel`` (csrc/kmeans.cu)."""
    xf = x.float().contiguous()
    n, d = xf.shape
    dev = xf.device
    ks = [int(k) for k in ks]
    nk, kmax = len(ks), max(ks)
    ks_t = torch.tensor(ks, dtype=torch.int32, device=dev)
    centers = torch.zeros(nk, nstart, kmax, d, device=dev)
    labels = torch.empty(nk, nstart, n, dtype=torch.int32, device=dev)
    mind = torch.empty(nk, nstart, n, device=dev)
    wss = torch.zeros(nk, nstart, kmax, device=dev)
    tot = torch.empty(nk, nstart, device=dev)
    iters = torch.empty(nk, nstart, dtype=torch.int32, device=dev)
    P = _native.ptr
    rc = _native.cuda().cml_kmeans_lloyd(P(xf), xf.stride(0), n, d, P(ks_t), nk, kmax, nstart, max_iter, ctypes.c_ulonglong(seed & (2**64 - 1)),
                                         P(centers), P(labels), P(mind), P(wss), P(tot), P(iters), _native.stream_ptr())
    _native.check(rc, "cml_kmeans_lloyd")
    best = tot.argmin(1).tolist()
    out = {}
    for e, k in enumerate(ks):
        b = best[e]
        out[k] = KMeansResult(centers[e, b, :k], labels[e, b].long(), float(tot[e, b]), wss[e, b, :k], int(iters[e, b]))
    return out


def _kmeans_gemm(x: torch.Tensor, k: int, nstart: int, max_iter: int, seed: int) -> KMeansResult:
    """Large cohorts (n x d beyond what one CTA should walk alone): Lloyd iterations built from the tcgen05
    squared-distance GEMM (``ops.gemm.sq_dists``), ``cml_kmeans_assign`` and ``cml_kmeans_update``, one restart at a time."""
    from ..ops import gemm as G
    xf = x.float().contiguous()
    n, d = xf.shape
    dev = xf.device
    gen = torch.Generator(device="cpu").manual_seed(seed + 7919 * k)
    P = _native.ptr
    lib = _native.cuda()
    best = None
    for _ in range(nstart):
        centers = xf[torch.randperm(n, generator=gen)[:k].to(dev)].clone()
        labels = torch.full((n,), -1, dtype=torch.int32, device=dev)
        new = torch.empty(n, dtype=torch.int32, device=dev)
        mind = torch.empty(n, device=dev)
        it = 0
        for it in range(1, max_iter + 1):
            d2 = G.sq_dists(xf, centers).contiguous()
            _native.check(lib.cml_kmeans_assign(P(d2), d2.stride(0), n, k, P(new), P(mind), _native.stream_ptr()), "cml_kmeans_assign")
            if torch.equal(new, labels):
                break
            labels, new = new, labels
            _native.check(lib.cml_kmeans_update(P(xf), xf.stride(0), n, d, k, P(labels), P(centers), None, _native.stream_ptr()),
                          "cml_kmeans_update")
        tot = float(mind.clamp_min(0).sum())
        if best is None or tot < best.tot_withinss:
            wss = torch.zeros(k, device=dev).index_add_(0, labels.long(), mind.clamp_min(0))
            best = KMeansResult(centers, labels.long(), tot, wss, it)
    return best


_FUSED_LIMIT = 1 << 22       # n * d above which one CTA per restart stops being the right shape


def kmeans(x: torch.Tensor, k: int, nstart: int = 25, max_iter: int = 100, seed: int = 1000) -> KMeansResult:
    if _native.on_device(x):
        n, d = x.shape
        if n * d <= _FUSED_LIMIT and 2 * k * d * 4 <= 200 * 1024:
            return _kmeans_device(x, [k], nstart, max_iter, seed)[k]
        return _kmeans_gemm(x, k, nstart, max_iter, seed)
    x = x.float()
    n, d = x.shape
    g = torch.Generator(device="cpu").manual_seed(seed + 7919 * k)
    # k-means++ seeding for every restart at once
    first = torch.randint(0, n, (nstart,), generator=g).to(x.device)
    centers = x[first][:, None, :].clone()                                    # [R, 1, d]
    for _ in range(1, k):
        d2 = _sqdist(x, centers).min(2).values                                # [R, n]
        prob = d2 / d2.sum(1, keepdim=True).clamp_min(1e-30)
        nxt = torch.multinomial(prob.cpu() + 1e-12, 1, generator=g).to(x.device).flatten()
        centers = torch.cat([centers, x[nxt][:, None, :]], 1)
    it = 0
    labels = None
    for it in range(1, max_iter + 1):
        d2 = _sqdist(x, centers)
        new = d2.argmin(2)                                                    # [R, n]
        if labels is not None and torch.equal(new, labels):
            break
        labels = new
        onehot = torch.nn.functional.one_hot(labels, k).float()               # [R, n, k]
        cnt = onehot.sum(1)                                                   # [R, k]
        sums = torch.einsum("rnk,nd->rkd", onehot, x)
        newc = sums / cnt.clamp_min(1)[:, :, None]
        centers = torch.where((cnt > 0)[:, :, None], newc, centers)          # an emptied cluster keeps its centre
    d2 = _sqdist(x, centers)
    mind = d2.gather(2, labels[:, :, None]).squeeze(2)
    tot = mind.sum(1)
    best = int(tot.argmin())
    onehot = torch.nn.functional.one_hot(labels[best], k).float()
    return KMeansResult(centers[best], labels[best], float(tot[best]), (onehot * mind[best][:, None]).sum(0), it)


def kmeans_sweep(x: torch.Tensor, ks: Sequence[int] = tuple(range(20, 1, -1)), nstart: int = 25, seed: int = 1000) -> Dict[int, KMeansResult]:
    """The reference's sweep k = 20, ..., 2.  On the kernel path all 19 x 25 problems are ONE launch."""
    n, d = x.shape
    if _native.on_device(x) and n * d <= _FUSED_LIMIT and 2 * max(ks) * d * 4 <= 200 * 1024:
        return _kmeans_device(x, list(ks), nstart, 100, seed)
    return {k: kmeans(x, k, nstart=nstart, seed=seed) for k in ks}
