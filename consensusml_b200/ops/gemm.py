"""tcgen05 GEMM front-end: Gram / RBF kernel matrices, squared distances, scaled products.

CUDA: ``cml_gemm_tn_bf16`` (csrc/gemm_tcgen05.cu - TMA -> smem ring -> tcgen05.mma -> TMEM -> fused
epilogue), operands rounded to bf16 by ``cml_to_bf16_rownorm`` which also emits the squared row norms the
distance epilogues need.  CPU: float64 torch reference.  Replaces the kernel-matrix evaluation inside
e1071::svm (cml_targetaml_seanalysis.Rmd:151-154), ``cor()`` (:895) and the proximity accumulation of
randomForest (:1037-1050)."""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import torch

from .. import _native

_I, _LL, _F, _P = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p
_native.register_signature("cml_to_bf16_rownorm", [_P, _LL, _P, _LL, _P, _I, _I, _F, _P, _P, _P])
_native.register_signature("cml_gemm_tn_bf16", [_I, _I, _I, _P, _LL, _P, _I, _I, _LL, _P, _LL, _I, _F, _P, _P, _P])

_native.register_signature("cml_gemm_set_variant", [_I])

EPI_SCALE, EPI_RBF, EPI_SQDIST = 0, 1, 2


_native.register_signature("cml_gemm_set_staged_grouping", [_I])


def set_staged_grouping(on: bool) -> None:
    """Grouped tile order inside each source block of the staged all-gather GEMM (opt-in until measured on hardware)."""
    _native.check(_native.cuda().cml_gemm_set_staged_grouping(int(bool(on))), "cml_gemm_set_staged_grouping")


def set_variant(variant: int) -> None:
    """0 = choose by shape, 1 = single-CTA tcgen05 kernel, 2 = CTA-pair (``cta_group::2``) kernel."""
    _native.check(_native.cuda().cml_gemm_set_variant(int(variant)), "cml_gemm_set_variant")


def _pad8(k: int) -> int:
    return (k + 7) // 8 * 8


def to_bf16_rows(x: torch.Tensor, *, shift: float = 0.0, col_shift: Optional[torch.Tensor] = None,
                 col_scale: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                 norm_out: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Row-major fp32 [rows, cols] -> (bf16 [rows, pad8(cols)] zero padded, squared norms of the rounded rows).
    Optional per-column centring / scaling is fused (z-scoring for e1071's ``scale = TRUE``)."""
    assert _native.on_device(x) and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    rows, cols = x.shape
    ld = _pad8(cols)
    y = out if out is not None else torch.empty(rows, ld, dtype=torch.bfloat16, device=x.device)
    nrm = norm_out if norm_out is not None else torch.empty(rows, dtype=torch.float32, device=x.device)
    P = _native.ptr
    rc = _native.cuda().cml_to_bf16_rownorm(P(x) if x.is_contiguous() else ctypes.c_void_p(x.data_ptr()), x.stride(0),
                                            P(y), ld, P(nrm), rows, cols, float(shift),
                                            P(col_shift.float().contiguous()) if col_shift is not None else None,
                                            P(col_scale.float().contiguous()) if col_scale is not None else None,
                                            _native.stream_ptr())
    _native.check(rc, "cml_to_bf16_rownorm")
    return y, nrm


def gemm_tn(a_bf16: torch.Tensor, b_sources: Sequence[torch.Tensor], n_total: int, *, epilogue: int = EPI_SCALE,
            alpha: float = 1.0, norm_a: Optional[torch.Tensor] = None,
            norm_b: Optional[Sequence[torch.Tensor]] = None, out: Optional[torch.Tensor] = None,
            b_ptrs: Optional[Sequence[int]] = None, norm_b_ptrs: Optional[Sequence[int]] = None,
            rows_per_src: Optional[int] = None, ldb: Optional[int] = None) -> torch.Tensor:
    """``C[M, n_total] = epilogue(A . B^T)`` with B given as equally sized row blocks ("sources").

    ``b_sources`` are bf16 [rows_per_src, ld] tensors (or pass raw device addresses via ``b_ptrs`` - that is
    how peer buffers of other GPUs enter for the fused all-gather)."""
    M, ld = a_bf16.shape
    K = ld
    if b_ptrs is None:
        b_ptrs = [t.data_ptr() for t in b_sources]
        rows_per_src = b_sources[0].shape[0]
        ldb = b_sources[0].shape[1]
    n_src = len(b_ptrs)
    ldc = (n_total + 3) // 4 * 4
    c = out if out is not None else torch.empty(M, ldc, dtype=torch.float32, device=a_bf16.device)
    arr_b = (ctypes.c_void_p * n_src)(*b_ptrs)
    if norm_b_ptrs is None and norm_b is not None:
        norm_b_ptrs = [t.data_ptr() for t in norm_b]
    arr_n = (ctypes.c_void_p * n_src)(*norm_b_ptrs) if norm_b_ptrs is not None else None
    rc = _native.cuda().cml_gemm_tn_bf16(M, n_total, K, _native.ptr(a_bf16), ld, arr_b, n_src, rows_per_src, ldb,
                                         _native.ptr(c), c.stride(0), epilogue, float(alpha),
                                         _native.ptr(norm_a) if norm_a is not None else None, arr_n,
                                         _native.stream_ptr())
    _native.check(rc, "cml_gemm_tn_bf16")
    return c[:, :n_total]


_native.register_signature("cml_gemm_tn_bf16_staged", [_I, _I, _I, _P, _LL, _P, _I, _I, _LL, _P, _LL, _I, _F, _P, _P, _P, _P, _P,
                                                         ctypes.c_uint, _I, _I, _P])
_native.register_signature("cml_gemm_staged_flags", [_I, _I], _I)


def gemm_tn_staged(a_bf16: torch.Tensor, peer_ptrs: Sequence[int], n_total: int, rows_per_src: int, ldb: int, stage: torch.Tensor,
                   stage_norm: Optional[torch.Tensor], flags: torch.Tensor, epoch: int, rank: int, *, epilogue: int = EPI_SCALE,
                   alpha: float = 1.0, norm_a: Optional[torch.Tensor] = None, peer_norm_ptrs: Optional[Sequence[int]] = None,
                   out: Optional[torch.Tensor] = None, n_copiers: int = 12) -> torch.Tensor:
    """Fused all-gather -> GEMM: ``peer_ptrs[s]`` is the device address of source block ``s`` (peer memory); copier CTAs
    of the same kernel stage the blocks into ``stage`` while the GEMM CTAs consume the tiles that have arrived."""
    M, K = a_bf16.shape
    n_src = len(peer_ptrs)
    n_copiers = max(1, min(n_copiers, _native.sm_count() // 2 - 1))      # the copiers must leave most of the machine to the GEMM
    ldc = (n_total + 3) // 4 * 4
    c = out if out is not None else torch.empty(M, ldc, dtype=torch.float32, device=a_bf16.device)
    arr_b = (ctypes.c_void_p * n_src)(*peer_ptrs)
    arr_n = (ctypes.c_void_p * n_src)(*peer_norm_ptrs) if peer_norm_ptrs is not None else None
    rc = _native.cuda().cml_gemm_tn_bf16_staged(M, n_total, K, _native.ptr(a_bf16), K, arr_b, n_src, rows_per_src, ldb,
                                                _native.ptr(c), c.stride(0), epilogue, float(alpha),
                                                _native.ptr(norm_a) if norm_a is not None else None, arr_n,
                                                _native.ptr(stage), _native.ptr(stage_norm) if stage_norm is not None else None,
                                                _native.ptr(flags), epoch & 0xFFFFFFFF, rank, n_copiers, _native.stream_ptr())
    _native.check(rc, "cml_gemm_tn_bf16_staged")
    return c[:, :n_total]


# Below one 64-wide K block the TMA -> tcgen05 pipeline has nothing to pipeline (and no such shape is part of the validated
# set, tests/test_gemm_gpu.py): feature dimensions that small go through plain torch on the same device.
MIN_K = 64


def rbf_kernel(x: torch.Tensor, z: Optional[torch.Tensor] = None, gamma: Optional[float] = None) -> torch.Tensor:
    """K(i, j) = exp(-gamma |x_i - z_j|^2); gamma defaults to 1/p as in e1071."""
    gamma = 1.0 / x.shape[1] if gamma is None else gamma
    if not _native.on_device(x) or x.shape[1] < MIN_K:
        zz = x if z is None else z
        d2 = torch.cdist(x.double(), zz.double()) ** 2
        return torch.exp(-gamma * d2).to(x.dtype)
    shift = float(x.mean())            # distances are shift-invariant; centring shrinks the norms that cancel
    xa, na = to_bf16_rows(x.float(), shift=shift)
    if z is None:
        xb, nb = xa, na
    else:
        xb, nb = to_bf16_rows(z.float(), shift=shift)
    return gemm_tn(xa, [xb], xb.shape[0], epilogue=EPI_RBF, alpha=gamma, norm_a=na, norm_b=[nb])


def linear_kernel(x: torch.Tensor, z: Optional[torch.Tensor] = None, scale: float = 1.0) -> torch.Tensor:
    if not _native.on_device(x) or x.shape[1] < MIN_K:
        zz = x if z is None else z
        return (scale * (x.double() @ zz.double().t())).to(x.dtype)
    xa, _ = to_bf16_rows(x.float())
    xb = xa if z is None else to_bf16_rows(z.float())[0]
    return gemm_tn(xa, [xb], xb.shape[0], epilogue=EPI_SCALE, alpha=scale)


def sq_dists(x: torch.Tensor, z: torch.Tensor) -> torch.Tensor:
    if not _native.on_device(x) or x.shape[1] < MIN_K:
        return (torch.cdist(x.double(), z.double()) ** 2).to(x.dtype)
    xa, na = to_bf16_rows(x.float())
    xb, nb = to_bf16_rows(z.float())
    return gemm_tn(xa, [xb], xb.shape[0], epilogue=EPI_SQDIST, norm_a=na, norm_b=[nb])
