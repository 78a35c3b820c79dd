"""Spearman / Pearson correlation matrices, hierarchical-clustering order, KDE, 2-D PCA - the numbers behind the
reference's figures (correlation heatmaps cml_targetaml_seanalysis.Rmd:883-971, density plots :973-1017,
ComplexHeatmap row/column ordering :536-613, ``fviz_cluster`` projection BuieRProj/KM_WHATEVER.Rmd:158).
Picture rendering itself is out of scope on this box (no graphics stack); the arrays are what ``report`` dumps.

CUDA: rows are rank-transformed and standardised by ``cml_standardize_rows`` and the p x p matrix is ONE tcgen05
GEMM (bf16 operands, fp32 accumulate).  ``exact=True`` keeps fp32 operands (library matmul) where bf16's
~3 significant digits matter.  CPU: float64 torch."""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import numpy as np
import torch

from .. import _native
from ..ops import gemm as G
from ..ops.metrics import average_ranks

_I, _LL, _F, _P = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p
_native.register_signature("cml_standardize_rows", [_P, _LL, _I, _I, _I, _P, _LL, _P, _P, _P])
_native.register_signature("cml_kde", [_P, _I, _F, _F, _F, _I, _P, _P])
_native.register_signature("cml_kmeans_assign", [_P, _LL, _I, _I, _P, _P, _P])


def _standardize(x: torch.Tensor, ranked: bool, cols: Optional[torch.Tensor], want_f32: bool):
    rows = x.shape[0]
    n = x.shape[1] if cols is None else int(cols.numel())
    ld = (n + 7) // 8 * 8
    zb = torch.empty(rows, ld, dtype=torch.bfloat16, device=x.device)
    zf = torch.empty(rows, n, dtype=torch.float32, device=x.device) if want_f32 else None
    ct = cols.to(device=x.device, dtype=torch.int32).contiguous() if cols is not None else None
    x = x.float().contiguous()
    rc = _native.cuda().cml_standardize_rows(_native.ptr(x), x.shape[1], rows, n, int(ranked), _native.ptr(zb), ld, _native.ptr(zf),
                                             _native.ptr(ct), _native.stream_ptr())
    _native.check(rc, "cml_standardize_rows")
    return zb, zf


def correlation(x: torch.Tensor, y: Optional[torch.Tensor] = None, method: str = "spearman",
                samples: Optional[Sequence[int]] = None, exact: bool = False) -> torch.Tensor:
    """Correlation between the rows of ``x`` (genes x samples) and the rows of ``y`` (default ``x``), optionally on a
    subset of sample columns (the reference repeats the heatmap on five sample subsets, SEA:889-969)."""
    ranked = method == "spearman"
    cols = torch.as_tensor(list(samples)) if samples is not None else None
    if not _native.on_device(x):
        def z(a):
            a = a.double()
            if cols is not None:
                a = a[:, cols]
            if ranked:
                a = torch.stack([average_ranks(r) for r in a])
            a = a - a.mean(1, keepdim=True)
            nrm = a.norm(dim=1, keepdim=True)
            return a / torch.where(nrm > 0, nrm, torch.ones_like(nrm))
        zx = z(x)
        zy = zx if y is None else z(y)
        return (zx @ zy.t()).to(torch.float32)
    zb, zf = _standardize(x, ranked, cols, exact)
    if y is None:
        yb, yf = zb, zf
    else:
        yb, yf = _standardize(y, ranked, cols, exact)
    if exact:
        return zf @ yf.t()
    return G.gemm_tn(zb, [yb], yb.shape[0]).clamp_(-1.0, 1.0)


HCLUST_METHODS = {"single": 0, "complete": 1, "average": 2, "weighted": 3, "ward": 4}


def hclust(dist: torch.Tensor, method: str = "complete"):
    """Agglomerative clustering of a precomputed distance matrix with the framework's own nearest-neighbour-chain
    implementation (csrc/host/hclust.cpp).  Returns (Z, leaves): the linkage matrix in scipy's convention ((n-1) x 4:
    ids, height, size; sorted by height) and the left-to-right leaf order.  scipy is only the test oracle."""
    lib = _native.host()
    d = dist.detach().double().cpu()
    d = (0.5 * (d + d.t())).clamp_min(0)
    d.fill_diagonal_(0.0)
    dn = np.ascontiguousarray(d.numpy())
    n = dn.shape[0]
    Z = np.zeros((max(n - 1, 0), 4))
    leaves = np.zeros(n, np.int32)
    lib.cml_host_hclust.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    rc = lib.cml_host_hclust(_native.np_ptr(dn), n, HCLUST_METHODS[method], _native.np_ptr(Z), _native.np_ptr(leaves))
    if rc != 0:
        raise ValueError(f"hclust: bad arguments (n={n}, method={method})")
    return Z, leaves


def hclust_order(dist: torch.Tensor, method: str = "complete") -> np.ndarray:
    """Leaf order of an agglomerative clustering of a distance matrix (ComplexHeatmap's default: complete linkage on
    Euclidean distances, cml_targetaml_seanalysis.Rmd:536-613).  The distance matrix comes from the tcgen05
    squared-distance GEMM (``ops.gemm.sq_dists``); the merges from ``hclust`` above."""
    return hclust(dist, method)[1].astype(np.int64)


def heatmap_order(x: torch.Tensor):
    """(row order, column order) for a clustered heatmap of ``x``: Euclidean distances between rows / columns."""
    dr = G.sq_dists(x.float(), x.float()).clamp_min(0).sqrt()
    xt = x.t().contiguous()
    dc = G.sq_dists(xt.float(), xt.float()).clamp_min(0).sqrt()
    return hclust_order(dr), hclust_order(dc)


def bw_nrd0(x: torch.Tensor) -> float:
    """R's default bandwidth ``bw.nrd0``: 0.9 min(sd, IQR/1.34) n^-1/5."""
    xd = x.double().flatten()
    n = xd.numel()
    sd = float(xd.std(unbiased=True))
    q = torch.quantile(xd, torch.tensor([0.25, 0.75], dtype=torch.float64, device=xd.device))
    iqr = float(q[1] - q[0]) / 1.34
    lo = min(sd, iqr) if iqr > 0 else sd
    if lo == 0:
        lo = abs(float(xd[0])) or 1.0
    return 0.9 * lo * n ** -0.2


def kde(x: torch.Tensor, n_grid: int = 512, bw: Optional[float] = None, cut: float = 3.0):
    """Gaussian KDE on a regular grid like R's ``density()`` (exact sum, not the binned FFT approximation).
    Returns (grid, density)."""
    xf = x.float().flatten().contiguous()
    bw = bw_nrd0(xf) if bw is None else bw
    lo, hi = float(xf.min()) - cut * bw, float(xf.max()) + cut * bw
    grid = torch.linspace(lo, hi, n_grid, device=x.device)
    if not _native.on_device(x):
        u = (grid[:, None].double() - xf[None, :].double()) / bw
        return grid, (torch.exp(-0.5 * u * u).sum(1) / (xf.numel() * bw * np.sqrt(2 * np.pi))).float()
    out = torch.empty(n_grid, dtype=torch.float32, device=x.device)
    rc = _native.cuda().cml_kde(_native.ptr(xf), xf.numel(), float(bw), lo, hi, n_grid, _native.ptr(out), _native.stream_ptr())
    _native.check(rc, "cml_kde")
    return grid, out


def pca2(x: torch.Tensor, k: int = 2):
    """First ``k`` principal-component scores of the rows of ``x`` (samples x features), like ``prcomp`` behind
    ``fviz_cluster``.  Uses the n x n Gram matrix (n << p here) and a symmetric eigen-decomposition.
    Returns (scores [n, k], explained variance ratio [k])."""
    xc = x.double() - x.double().mean(0, keepdim=True)
    gram = xc @ xc.t()
    evals, evecs = torch.linalg.eigh(gram)
    idx = torch.argsort(evals, descending=True)[:k]
    lam = evals[idx].clamp_min(0)
    scores = evecs[:, idx] * lam.sqrt()[None, :]
    # deterministic sign: largest-magnitude score of each component positive
    sign = torch.sign(scores[scores.abs().argmax(0), torch.arange(k)])
    return (scores * sign[None, :]).float(), (lam / evals.clamp_min(0).sum()).float()
