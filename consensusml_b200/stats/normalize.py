"""Count normalisation: library sizes, CPM expression filter, TMM factors, log2-CPM (K1-K3).

API mirrors the edgeR sequence the reference runs three times
(JSmith_code/Normalization_and_Batch_Effect_Investigation.Rmd:49-63 pan-cancer;
JSmith_code/Differential_Expression_and_Lasso.Rmd:217-227 AML; seobjects/make_seobj_targetaml.R:88-96).
``counts`` is an int32 genes x samples tensor.  CUDA tensors run the kernels of csrc/norm.cu; CPU tensors
run the float64 oracle (consensusml_b200/oracle/limma.py)."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from .. import _native
from ..oracle import limma as O

_I, _LL, _D, _P = ctypes.c_int, ctypes.c_longlong, ctypes.c_double, ctypes.c_void_p
_native.register_signature("cml_lib_sizes", [_P, _I, _I, _P, _P])
_native.register_signature("cml_cpm_pass_count", [_P, _I, _I, _P, _D, _P, _P])
_native.register_signature("cml_nonzero_rows", [_P, _I, _I, _P, _P])
_native.register_signature("cml_quantile75", [_P, _I, _I, _P, _I, _P, _P, _P])
_native.register_signature("cml_tmm_factors", [_P, _I, _I, _P, _P, _I, _D, _D, _D, _P, _P])
_native.register_signature("cml_logcpm", [_P, _I, _I, _P, _I, _P, _D, _D, _P, _P, _LL, _P])


def _as_counts(counts: torch.Tensor) -> torch.Tensor:
    if counts.dtype != torch.int32:
        counts = counts.to(torch.int32)
    return counts.contiguous()


def lib_sizes(counts: torch.Tensor) -> torch.Tensor:
    """Column sums (``DGEList`` library sizes), float64 [S]."""
    counts = _as_counts(counts)
    if not _native.on_device(counts):
        return counts.to(torch.float64).sum(0)
    G, S = counts.shape
    out = torch.empty(S, dtype=torch.int64, device=counts.device)
    rc = _native.cuda().cml_lib_sizes(_native.ptr(counts), G, S, _native.ptr(out), _native.stream_ptr())
    _native.check(rc, "cml_lib_sizes")
    return out.to(torch.float64)


def cpm_pass_counts(counts: torch.Tensor, lib: torch.Tensor, min_cpm: float = 1.0) -> torch.Tensor:
    """Per gene, the number of samples whose CPM is at least ``min_cpm`` (int32 [G]).  When the cohort is
    sample-sharded these per-gene counts are what gets all-reduced across ranks."""
    counts = _as_counts(counts)
    if not _native.on_device(counts):
        return torch.from_numpy((O.cpm(counts.numpy().astype(np.float64), lib.numpy()) >= min_cpm).sum(1).astype(np.int32))
    G, S = counts.shape
    out = torch.empty(G, dtype=torch.int32, device=counts.device)
    rc = _native.cuda().cml_cpm_pass_count(_native.ptr(counts), G, S, _native.ptr(lib.double().contiguous()),
                                           float(min_cpm), _native.ptr(out), _native.stream_ptr())
    _native.check(rc, "cml_cpm_pass_count")
    return out


def cpm_filter(counts: torch.Tensor, min_cpm: float = 1.0, min_samples: float = 5, strict: bool = False,
               lib: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Boolean gene mask.  The reference's three threshold rules:
    ``>= max(2, 0.05 n)`` (Normalization...Rmd:55), ``>= 5`` (Differential...Rmd:221), ``> 0.05 n`` (voom_DE, strict)."""
    lib = lib_sizes(counts) if lib is None else lib
    n_ok = cpm_pass_counts(counts, lib, min_cpm)
    return n_ok > min_samples if strict else n_ok >= min_samples


def upper_quartile_factors(counts: torch.Tensor, lib: torch.Tensor, all_rows: bool = False) -> torch.Tensor:
    """edgeR's ``f75`` per sample: the 75 % quantile of count / library size over the genes with a count anywhere (the
    statistic TMM picks its reference sample with).  float64 [S].  ``all_rows``: take every row (a sample shard cannot tell
    from its own columns whether a gene has counts elsewhere; rows that passed the CPM filter all do)."""
    counts = _as_counts(counts)
    lib = lib.double().contiguous()
    G, S = counts.shape
    if not _native.on_device(counts):
        c = counts.numpy().astype(np.float64)
        if not all_rows:
            c = c[c.sum(1) > 0]
        return torch.from_numpy(np.quantile(c / lib.numpy()[None, :], 0.75, axis=0))
    lb = _native.cuda()
    P = _native.ptr
    mask = torch.empty(G, dtype=torch.uint8, device=counts.device)
    if all_rows:
        mask.fill_(1)
    else:
        _native.check(lb.cml_nonzero_rows(P(counts), G, S, P(mask), _native.stream_ptr()), "cml_nonzero_rows")
    f75 = torch.empty(S, dtype=torch.float64, device=counts.device)
    _native.check(lb.cml_quantile75(P(counts), G, S, P(mask), int(mask.sum().item()), P(lib), P(f75), _native.stream_ptr()), "cml_quantile75")
    return f75


def tmm_factors(counts: torch.Tensor, lib: Optional[torch.Tensor] = None, *, logratio_trim: float = 0.3,
                sum_trim: float = 0.05, a_cutoff: float = -1e10, ref_column: Optional[int] = None, normalize: bool = True) -> torch.Tensor:
    """edgeR ``calcNormFactors(method="TMM")`` on a (filtered) count matrix; float64 [S] with geometric mean 1
    (``normalize=False`` returns the raw factors - the sample-sharded path normalises over all ranks' samples)."""
    counts = _as_counts(counts)
    lib = lib_sizes(counts) if lib is None else lib.double()
    if not _native.on_device(counts):
        return torch.from_numpy(O.tmm_factors(counts.numpy().astype(np.float64), lib.numpy(), ref_column, normalize=normalize))
    lib = lib.contiguous()
    G, S = counts.shape
    lb = _native.cuda()
    P = _native.ptr
    mask = torch.empty(G, dtype=torch.uint8, device=counts.device)
    _native.check(lb.cml_nonzero_rows(P(counts), G, S, P(mask), _native.stream_ptr()), "cml_nonzero_rows")
    n_rows = int(mask.sum().item())
    if ref_column is None:
        f75 = torch.empty(S, dtype=torch.float64, device=counts.device)
        _native.check(lb.cml_quantile75(P(counts), G, S, P(mask), n_rows, P(lib), P(f75), _native.stream_ptr()),
                      "cml_quantile75")
        if float(f75.median()) < 1e-20:
            ref_column = int(torch.argmax(counts.double().sqrt().sum(0)).item())
        else:
            ref_column = int(torch.argmin((f75 - f75.mean()).abs()).item())
    f = torch.empty(S, dtype=torch.float64, device=counts.device)
    _native.check(lb.cml_tmm_factors(P(counts), G, S, P(mask), P(lib), int(ref_column), logratio_trim, sum_trim,
                                     a_cutoff, P(f), _native.stream_ptr()), "cml_tmm_factors")
    return f / torch.exp(torch.log(f).mean()) if normalize else f


def log_cpm(counts: torch.Tensor, lib: torch.Tensor, norm_factors: Optional[torch.Tensor] = None,
            prior_count: float = 1.0, rows: Optional[torch.Tensor] = None, sample_major: bool = False,
            mean_eff: Optional[float] = None) -> torch.Tensor:
    """``cpm(dge, log=TRUE, prior.count=p)`` -> fp32.  ``rows`` selects genes (int indices);
    ``sample_major=True`` writes the transposed samples x genes model matrix directly.  ``mean_eff`` overrides the mean
    effective library size that scales the prior count (a sample shard passes the mean over ALL samples)."""
    counts = _as_counts(counts)
    eff = lib.double() * (1.0 if norm_factors is None else norm_factors.double())
    if not _native.on_device(counts):
        c = counts.numpy().astype(np.float64)
        if rows is not None:
            c = c[rows.numpy()]
        out = torch.from_numpy(O.log_cpm(c, eff.numpy(), None, prior_count, mean_lib=mean_eff).astype(np.float32))
        return out.t().contiguous() if sample_major else out
    G, S = counts.shape
    n_rows = G if rows is None else int(rows.numel())
    rows_t = rows.to(device=counts.device, dtype=torch.int32).contiguous() if rows is not None else None
    eff = eff.contiguous()
    out = out_t = None
    ld_t = 0
    if sample_major:
        out_t = torch.empty(S, n_rows, dtype=torch.float32, device=counts.device)
        ld_t = n_rows
    else:
        out = torch.empty(n_rows, S, dtype=torch.float32, device=counts.device)
    rc = _native.cuda().cml_logcpm(_native.ptr(counts), G, S, _native.ptr(rows_t), n_rows, _native.ptr(eff),
                                   float(prior_count), float(eff.mean()) if mean_eff is None else float(mean_eff), _native.ptr(out),
                                   _native.ptr(out_t), ld_t,
                                   _native.stream_ptr())
    _native.check(rc, "cml_logcpm")
    return out_t if sample_major else out


@dataclass
class Normalized:
    keep: torch.Tensor          # bool [G]
    lib_size: torch.Tensor      # float64 [S] (original library sizes, as DGEList subsetting keeps them)
    norm_factors: torch.Tensor  # float64 [S]
    logcpm: torch.Tensor        # fp32 [G_kept, S]


def tmm_normalize(counts: torch.Tensor, min_cpm: float = 1.0, min_samples: float = 5, strict: bool = False,
                  prior_count: float = 1.0) -> Normalized:
    """The whole block ``DGEList -> filter -> calcNormFactors -> cpm(log=TRUE)``
    (Differential_Expression_and_Lasso.Rmd:217-227 with ``min_samples=5``)."""
    counts = _as_counts(counts)
    lib = lib_sizes(counts)
    keep = cpm_filter(counts, min_cpm, min_samples, strict, lib)
    rows = torch.nonzero(keep).flatten()
    sub = counts[rows].contiguous()
    nf = tmm_factors(sub, lib)
    lc = log_cpm(sub, lib, nf, prior_count)
    return Normalized(keep, lib, nf, lc)
