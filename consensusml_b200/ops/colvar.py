"""Per-gene variance and top-k selection (K12).

Replaces ``ldf.var(axis=0)`` + ``sorted(...)[:1000]`` (scripts/model_comp.py:62-79) and
``order(rowVars(...))[1:100]`` (VikasP/AML.R:157-158).  On CUDA the reduction is ``column_moments_kernel`` (one
Welford pass per column over the rows a mask selects, Chan merges in a fixed order) and the selection is
``topk_select_kernel`` (radix select, ties to the lower index) - csrc/select.cu.  The (count, mean, M2) triple is the
mergeable summary: a sample-sharded cohort all-reduces it with ``merge_moments`` / ``sharded_moments`` before the top-k."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from .. import _native

_I, _LL, _P, _SZ = ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_size_t
_native.register_signature("cml_column_moments_scratch_bytes", [_I, _I], _SZ)
_native.register_signature("cml_column_moments", [_P, _LL, _I, _I, _P, _P, _P, _P])
_native.register_signature("cml_topk_select", [_P, _I, _I, _I, _P, _P, _P])


def column_moments(x: torch.Tensor, row_mask: Optional[torch.Tensor] = None):
    """(count, mean, M2) per column of a samples x genes matrix over the rows ``row_mask`` selects, in float64."""
    n, p = x.shape
    if not _native.on_device(x):
        xd = x.double() if row_mask is None else x[row_mask.bool()].double()
        cnt = xd.shape[0]
        mean = xd.mean(0) if cnt else torch.zeros(p, dtype=torch.float64)
        return cnt, mean, ((xd - mean) ** 2).sum(0)
    xf = x.float()
    if xf.stride(1) != 1:
        xf = xf.contiguous()
    lib = _native.cuda()
    scratch = torch.empty(lib.cml_column_moments_scratch_bytes(n, p), dtype=torch.uint8, device=x.device)
    out = torch.empty(3, p, dtype=torch.float64, device=x.device)
    mask = row_mask.to(device=x.device, dtype=torch.uint8).contiguous() if row_mask is not None else None
    rc = lib.cml_column_moments(_native.ptr(xf), xf.stride(0), n, p, _native.ptr(mask), _native.ptr(scratch), _native.ptr(out),
                                _native.stream_ptr())
    _native.check(rc, "cml_column_moments")
    cnt = int(out[0, 0].item()) if p else 0
    return cnt, out[1], out[2]


def merge_moments(a, b):
    """Chan et al. pairwise merge of two (count, mean, M2) summaries."""
    na, ma, sa = a
    nb, mb, sb = b
    n = na + nb
    if n == 0:
        return a
    d = mb - ma
    return n, ma + d * (nb / n), sa + sb + d * d * (na * nb / n)


def sharded_moments(x_local: torch.Tensor, row_mask: Optional[torch.Tensor] = None):
    """Moments of a sample-sharded cohort: local kernel pass, then one all-gather of the (count, mean, M2) triples
    (3 p numbers per rank) merged in rank order - the same result on every rank."""
    import torch.distributed as dist
    local = column_moments(x_local, row_mask)
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return local
    p = x_local.shape[1]
    pack = torch.empty(2 * p + 1, dtype=torch.float64, device=x_local.device)
    pack[0] = local[0]
    pack[1:p + 1], pack[p + 1:] = local[1], local[2]
    parts = [torch.empty_like(pack) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, pack)
    acc = (int(parts[0][0].item()), parts[0][1:p + 1], parts[0][p + 1:])
    for t in parts[1:]:
        acc = merge_moments(acc, (int(t[0].item()), t[1:p + 1], t[p + 1:]))
    return acc


def column_variance(x: torch.Tensor, row_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    n, _, m2 = column_moments(x, row_mask)
    return m2 / max(n - 1, 1)


def topk(values: torch.Tensor, k: int, largest: bool = True) -> torch.Tensor:
    """Indices of the k largest (smallest) values, best first, ties by index - the order of a stable sort."""
    p = values.numel()
    k = min(k, p)
    if not _native.on_device(values):
        return torch.argsort(values, descending=largest, stable=True)[:k]
    v = values.float().contiguous()
    idx = torch.empty(k, dtype=torch.int32, device=v.device)
    cnt = torch.zeros(1, dtype=torch.int32, device=v.device)
    rc = _native.cuda().cml_topk_select(_native.ptr(v), p, k, int(largest), _native.ptr(idx), _native.ptr(cnt), _native.stream_ptr())
    _native.check(rc, "cml_topk_select")
    idx = idx.long()
    # the kernel returns the set (better-than-threshold in index order, then ties); order those k by value
    order = torch.argsort(v[idx], descending=largest, stable=True)
    return idx[order]


def variance_topk(x: torch.Tensor, k: int, largest: bool = True, row_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Indices of the k columns with the largest (or smallest) sample variance, ordered like the reference's
    ``sorted(..., reverse=...)`` (stable, ties by column index)."""
    return topk(column_variance(x, row_mask), k, largest)
