"""Normalisation and differential expression for a SAMPLE-SHARDED count matrix (the 60,488 x 474 pan-cancer ingest of
JSmith_code/Normalization_and_Batch_Effect_Investigation.Rmd:49-66 and its limma-voom pass, spread over the ranks).

Every rank holds all genes of its block of samples (``counts_local`` [G, S_local], samples partitioned by
``parallel.dist.shard_bounds``).  The collectives are the ones SURVEY.md K1-K5 names:

  K1  CPM filter      per-gene pass counts are all-reduced (G ints); library sizes are all-gathered (S floats)
  K2  TMM             per-sample work stays local; the reference sample is chosen from the gathered upper quartiles, its
                      column is broadcast by its owner (G ints), the log-factors' mean is all-reduced (geometric mean 1)
  K3  log2-CPM        elementwise, local, with the GLOBAL mean effective library size
  K4 / K5 voom, lmFit, eBayes   are per-gene: the kept counts go through the sample-shard -> gene-shard transpose
                      (``parallel.transpose``: the fused all-to-all kernel on CUDA) and ``stats.de.voom_core`` runs
                      gene-sharded - local kernels on the rank's gene block, five all-gathers of G-length vectors for the
                      steps that need every gene (LOWESS trend, eBayes hyper-parameters, mixture prior, BH).

Results are the same on every rank and equal to the single-process functions up to floating-point summation order."""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from ..parallel.dist import shard_bounds
from . import de as DE
from . import normalize as N


def _wr():
    return (dist.get_world_size(), dist.get_rank()) if dist.is_initialized() else (1, 0)


def gather_var(t: torch.Tensor) -> torch.Tensor:
    """Concatenate a 1-D tensor over the ranks (rank order); lengths may differ."""
    world, _ = _wr()
    if world == 1:
        return t
    n = torch.tensor([t.numel()], dtype=torch.int64, device=t.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes)
    pad = torch.zeros(m, dtype=t.dtype, device=t.device)
    pad[: t.numel()] = t
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)])


def tmm_normalize_sharded(counts_local: torch.Tensor, min_cpm: float = 1.0, min_samples: float = 5, strict: bool = False,
                          prior_count: float = 1.0, want_logcpm: bool = True) -> dict:
    """Sample-sharded ``DGEList -> filter -> calcNormFactors -> cpm(log = TRUE)``.  Returns ``keep`` [G] bool,
    ``lib_size`` / ``norm_factors`` [S_total] float64 (global, rank order) and ``logcpm`` [G_kept, S_local] fp32."""
    world, rank = _wr()
    counts_local = counts_local.to(torch.int32).contiguous()
    dev = counts_local.device
    lib_local = N.lib_sizes(counts_local)
    lib = gather_var(lib_local)
    n_ok = N.cpm_pass_counts(counts_local, lib_local, min_cpm).to(torch.int64)
    if world > 1:
        dist.all_reduce(n_ok)                                       # K1: one all-reduce of G integers
    keep = n_ok > min_samples if strict else n_ok >= min_samples
    rows = torch.nonzero(keep).flatten()
    sub = counts_local[rows].contiguous()
    # ---- TMM: choose the reference sample globally, fetch its column, then everything is per local sample
    s_local = sub.shape[1]
    f75_local = _upper_quartiles(sub, lib_local)
    f75 = gather_var(f75_local)
    if float(f75.median()) < 1e-20:
        score = gather_var(sub.double().sqrt().sum(0))
        ref = int(torch.argmax(score))
    else:
        ref = int(torch.argmin((f75 - f75.mean()).abs()))
    S = lib.numel()
    owner = next(r for r in range(world) if shard_bounds(S, world, r)[0] <= ref < shard_bounds(S, world, r)[1])
    ref_col = torch.empty(sub.shape[0], dtype=torch.int32, device=dev)
    if rank == owner:
        ref_col.copy_(sub[:, ref - shard_bounds(S, world, owner)[0]])
    if world > 1:
        dist.broadcast(ref_col, src=owner)                          # K2: the reference column, G_kept integers
    ext = torch.cat([sub, ref_col[:, None]], 1).contiguous()
    lib_ext = torch.cat([lib_local, lib[ref:ref + 1]])
    f_raw = _tmm_raw(ext, lib_ext, s_local)[:s_local]
    logsum = torch.log(f_raw).sum().reshape(1)
    if world > 1:
        dist.all_reduce(logsum)                                     # K2: geometric mean 1 over ALL samples
    nf_local = f_raw / torch.exp(logsum / S)
    nf = gather_var(nf_local)
    out = {"keep": keep, "rows": rows, "lib_size": lib, "norm_factors": nf, "lib_local": lib_local, "nf_local": nf_local, "sub": sub}
    if want_logcpm:
        eff = lib * nf
        out["logcpm"] = N.log_cpm(sub, lib_local, nf_local, prior_count, mean_eff=float(eff.mean()))      # K3: global mean library size
    return out


def _upper_quartiles(sub: torch.Tensor, lib_local: torch.Tensor) -> torch.Tensor:
    """edgeR's ``f75`` of every local sample (75 % quantile of count / libsize over the genes with any count)."""
    return N.upper_quartile_factors(sub, lib_local, all_rows=True)


def _tmm_raw(ext: torch.Tensor, lib_ext: torch.Tensor, ref_index: int) -> torch.Tensor:
    """Un-normalised TMM factors of every column of ``ext`` against column ``ref_index``."""
    return N.tmm_factors(ext, lib_ext, ref_column=ref_index, normalize=False)


def voom_de_sharded(counts_local: torch.Tensor, group_all: np.ndarray, *, min_frac: float = 0.05, p_value: float = 0.05,
                    lfc: float = 1.0, span: float = 0.5, design: Optional[np.ndarray] = None, contrast: Optional[np.ndarray] = None):
    """limma-voom DE of a sample-sharded cohort.  ``group_all`` (or ``design``) covers ALL samples in global order.
    Returns the same ``DEResult`` as ``stats.de.voom_de`` (full-length vectors, identical on every rank)."""
    from ..parallel import transpose as TR
    world, rank = _wr()
    group_all = np.asarray(group_all)
    S = group_all.shape[0] if design is None else np.asarray(design).shape[0]
    nrm = tmm_normalize_sharded(counts_local, 1.0, min_frac * S, strict=True, want_logcpm=False)
    sub = nrm["sub"]                                                   # [G_kept, S_local]
    lib_eff = (nrm["lib_size"] * nrm["norm_factors"]).contiguous()
    # sample shards -> gene shards (path 4): this rank's block of kept genes, all samples
    genes = TR.to_gene_shards(sub.t().contiguous(), S)                 # [G_owned, S]
    X = DE._design_two_group(group_all) if design is None else np.asarray(design, np.float64)
    cvec = np.array([1.0, -1.0]) if contrast is None else np.asarray(contrast, np.float64)
    if not DE._native.on_device(genes):
        raise NotImplementedError("gene-sharded voom runs on the kernel path (CUDA, or CPU tensors inside _native.emulation())")
    r = DE.voom_core(genes.contiguous(), lib_eff, X, cvec, span=span, p_value=p_value, lfc=lfc, gather=gather_var)
    c = lambda v: v.cpu().numpy()
    return DE.DEResult(nrm["keep"], c(nrm["rows"]), nrm["lib_size"], nrm["norm_factors"], c(r["coef_c"]), c(r["amean"]), c(r["t"]), c(r["p"]),
                       c(r["adj"]), c(r["lods"]), r["s20"], r["df0"], r["var_prior"], c(r["idx"]), tuple(c(v) for v in r["trend"]))
