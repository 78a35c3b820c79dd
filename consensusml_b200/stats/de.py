"""limma-voom differential expression prefilter (K4-K5) and the DEG table writer.

``voom_de`` restates the reference's ``voom_DE(counts.df, ref, pheno)``
(JSmith_code/Limma_Voom_DE_Function.R:9-49; called at Differential_Expression_and_Lasso.Rmd:254 and :285):
CPM filter (> 5 % of samples), TMM, ``~0+group`` design, voom, per-gene WLS, contrast non-ref minus ref,
eBayes, ``topTable(p.value=0.05, adjust="BH", lfc=1, sort.by="P")``.  On CUDA every per-gene stage is a
kernel of csrc/voom.cu (one warp per gene); the scalar hyper-parameter fits (squeezeVar prior, B-statistic
prior variance) are a handful of reductions finished on the host - when genes are sharded across ranks
those are the two scalar all-reduces.  On CPU tensors the float64 oracle runs instead.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np
import pandas as pd
import torch
from scipy import special

from .. import _native
from ..oracle import limma as O
from . import normalize as N

_I, _D, _P = ctypes.c_int, ctypes.c_double, ctypes.c_void_p
_native.register_signature("cml_voom_ols", [_P, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P])
_native.register_signature("cml_lowess_points", [_P, _P, _P, _I, _I, _P, _I, _P, _P])
_native.register_signature("cml_voom_wls", [_P, _I, _I, _I, _P, _P, _P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P])
_native.register_signature("cml_ebayes", [_P, _P, _P, _P, _D, _I, _D, _D, _D, _D, _D, _P, _P, _P, _P, _P, _P])


@dataclass
class DEResult:
    keep: torch.Tensor              # bool [G_all] genes that passed the CPM filter
    gene_index: np.ndarray          # indices (into the full gene list) of the kept genes
    lib_size: torch.Tensor
    norm_factors: torch.Tensor
    logFC: np.ndarray               # per kept gene
    AveExpr: np.ndarray
    t: np.ndarray
    P_Value: np.ndarray
    adj_P_Val: np.ndarray
    B: np.ndarray
    s2_prior: float
    df_prior: float
    var_prior: float
    deg_order: np.ndarray           # kept-gene indices of the DEGs, sorted by P
    trend: Optional[tuple] = None

    def table(self, gene_ids: Sequence[str]) -> pd.DataFrame:
        """limma ``topTable`` frame (index = gene id; columns logFC, AveExpr, t, P.Value, adj.P.Val, B)."""
        i = self.deg_order
        ids = [gene_ids[self.gene_index[j]] for j in i]
        return pd.DataFrame({"logFC": self.logFC[i], "AveExpr": self.AveExpr[i], "t": self.t[i],
                             "P.Value": self.P_Value[i], "adj.P.Val": self.adj_P_Val[i], "B": self.B[i]}, index=ids)


# ------------------------------------------------------------------------------------------- lowess on device
def lowess_cuda(x: torch.Tensor, y: torch.Tensor, f: float = 0.5, iters: int = 3) -> torch.Tensor:
    """Robust lowess of sorted (x, y) float64 CUDA vectors; local fits run in ``cml_lowess_points``."""
    n = x.numel()
    xh = x.cpu().numpy()
    delta = 0.01 * (xh[-1] - xh[0])
    pts_np = O.lowess_eval_points(xh, delta)
    pts = torch.as_tensor(pts_np, dtype=torch.int32, device=x.device)
    m = pts.numel()
    ns = int(max(min(int(f * n + 1e-7), n), 2))
    lib = _native.cuda()
    P = _native.ptr
    fit_pts = torch.empty(m, dtype=torch.float64, device=x.device)
    xp = x[pts.long()]
    rw = None
    ys = None
    for it in range(iters + 1):
        rc = lib.cml_lowess_points(P(x), P(y), P(rw), n, ns, P(pts), m, P(fit_pts), _native.stream_ptr())
        _native.check(rc, "cml_lowess_points")
        # linear interpolation between evaluated points; tied x inherit the value of their evaluated twin
        seg = torch.searchsorted(xp, x, right=True).clamp(1, m - 1) if m > 1 else torch.zeros(n, dtype=torch.long, device=x.device)
        if m > 1:
            x0, x1 = xp[seg - 1], xp[seg]
            y0, y1 = fit_pts[seg - 1], fit_pts[seg]
            alpha = torch.where(x1 > x0, (x - x0) / (x1 - x0), torch.zeros_like(x))
            ys = alpha * y1 + (1 - alpha) * y0
        else:
            ys = fit_pts.expand(n).clone()
        if it >= iters:
            break
        res = (y - ys).abs()
        sc = res.sum() / n
        srt = torch.sort(res).values
        m1 = n // 2
        cmad = 3.0 * (srt[m1] + srt[n - m1 - 1]) if n % 2 == 0 else 6.0 * srt[m1]
        if float(cmad) < 1e-7 * float(sc):
            break
        c9, c1 = 0.999 * cmad, 0.001 * cmad
        rw = torch.where(res <= c1, torch.ones_like(res),
                         torch.where(res <= c9, (1 - (res / cmad) ** 2) ** 2, torch.zeros_like(res))).contiguous()
    return ys


# ------------------------------------------------------------------------------------------- voom DE
def _design_two_group(group: np.ndarray) -> np.ndarray:
    return np.stack([group == 1, group == 0], 1).astype(np.float64)


def voom_de(counts: torch.Tensor, group, *, min_frac: float = 0.05, p_value: float = 0.05, lfc: float = 1.0,
            span: float = 0.5, proportion: float = 0.01, stdev_coef_lim=(0.1, 4.0), design: Optional[np.ndarray] = None,
            contrast: Optional[np.ndarray] = None) -> DEResult:
    """Two-group limma-voom DE on raw counts (genes x samples).  ``group`` = 1 for the non-reference level.
    A custom ``design`` ([S, k<=8]) and ``contrast`` ([k]) may replace the two-group default."""
    group = np.asarray(group)
    if not _native.on_device(counts):
        c = counts.numpy().astype(np.float64)
        r = O.voom_de(c, group, min_frac) if design is None else _oracle_general(c, design, contrast, min_frac)
        eb = r["ebayes"]
        adj = O.bh_adjust(eb.p_value)
        return DEResult(torch.from_numpy(r["keep"]), np.flatnonzero(r["keep"]), torch.from_numpy(r["lib_size"]),
                        torch.from_numpy(r["norm_factors"]), r["coef"], r["fit"].amean, eb.t, eb.p_value, adj, eb.lods,
                        eb.s2_prior, eb.df_prior, eb.var_prior, r["deg_index"], r["trend"])
    counts = counts.to(torch.int32).contiguous()
    G_all, S = counts.shape
    lib = N.lib_sizes(counts)
    keep = N.cpm_filter(counts, 1.0, min_frac * S, strict=True, lib=lib)
    rows = torch.nonzero(keep).flatten()
    sub = counts[rows].contiguous()
    nf = N.tmm_factors(sub, lib)
    lib_eff = (lib * nf).contiguous()
    X = _design_two_group(group) if design is None else np.asarray(design, np.float64)
    cvec = np.array([1.0, -1.0]) if contrast is None else np.asarray(contrast, np.float64)
    r = voom_core(sub, lib_eff, X, cvec, span=span, proportion=proportion, stdev_coef_lim=stdev_coef_lim, p_value=p_value, lfc=lfc)
    c = lambda v: v.cpu().numpy()
    return DEResult(keep, c(rows), lib, nf, c(r["coef_c"]), c(r["amean"]), c(r["t"]), c(r["p"]), c(r["adj"]), c(r["lods"]), r["s20"], r["df0"],
                    r["var_prior"], c(r["idx"]), tuple(c(v) for v in r["trend"]))


def voom_core(sub: torch.Tensor, lib_eff: torch.Tensor, X: np.ndarray, cvec: np.ndarray, *, span: float = 0.5, proportion: float = 0.01,
              stdev_coef_lim=(0.1, 4.0), p_value: float = 0.05, lfc: float = 1.0, gather=None) -> dict:
    """voom weights -> weighted least squares -> contrast -> eBayes -> BH on the kernel path, for the gene rows in ``sub``
    ([G_local, S] int32 counts of ALL samples, already filtered).

    ``gather`` makes it gene-sharded (stats/sharded.py): every rank passes its block of genes and a function that
    concatenates a per-gene vector over the ranks in gene order.  The per-gene kernels (voom_ols, voom_wls, ebayes) then run
    on the local block only; the few steps that need every gene - the LOWESS mean-variance trend, the eBayes hyper-parameters,
    the top-|t| mixture fit and the BH adjustment - run on the gathered G-length vectors, identically on every rank.  The
    returned vectors are full length."""
    dev = sub.device
    ag = gather if gather is not None else (lambda v: v)
    Gl, S = sub.shape
    k = X.shape[1]
    hat = np.linalg.solve(X.T @ X, X.T)
    Xd = torch.as_tensor(X, device=dev).contiguous()
    Hd = torch.as_tensor(hat, device=dev).contiguous()
    cd = torch.as_tensor(cvec, device=dev).contiguous()
    f64 = dict(dtype=torch.float64, device=dev)
    coef_ols = torch.empty(Gl, k, **f64)
    sig_ols = torch.empty(Gl, **f64)
    amean_l = torch.empty(Gl, **f64)
    lb = _native.cuda()
    P = _native.ptr
    st = _native.stream_ptr()
    lib_eff = lib_eff.contiguous()
    if Gl:
        _native.check(lb.cml_voom_ols(P(sub), Gl, S, k, P(lib_eff), P(Xd), P(Hd), P(coef_ols), P(sig_ols), P(amean_l), None, st),
                      "cml_voom_ols")
    # mean-variance trend (needs the points of every gene)
    amean = ag(amean_l)
    G = amean.numel()
    sx = amean + torch.log2(lib_eff + 1.0).mean() - np.log2(1e6)
    sy = ag(sig_ols.sqrt())
    nz = ag((sub.sum(1) > 0).to(torch.uint8)).bool()
    sxk, syk = sx[nz], sy[nz]
    order = torch.argsort(sxk, stable=True)
    lx, ly_in = sxk[order].contiguous(), syk[order].contiguous()
    ly = lowess_cuda(lx, ly_in, f=span)
    ux, inv = torch.unique_consecutive(lx, return_inverse=True)
    uy = torch.zeros_like(ux).index_add_(0, inv, ly) / torch.zeros_like(ux).index_add_(0, inv, torch.ones_like(ly))
    ux, uy = ux.contiguous(), uy.contiguous()
    coef = torch.empty(Gl, k, **f64)
    su = torch.empty(Gl, k, **f64)
    sigma_l = torch.empty(Gl, **f64)
    coef_c_l = torch.empty(Gl, **f64)
    su_c_l = torch.empty(Gl, **f64)
    if Gl:
        _native.check(lb.cml_voom_wls(P(sub), Gl, S, k, P(lib_eff), P(Xd), P(coef_ols), P(ux), P(uy), ux.numel(), P(cd),
                                      P(coef), P(su), P(sigma_l), P(coef_c_l), P(su_c_l), None, st), "cml_voom_wls")
    # ---- empirical Bayes hyper-parameters: reductions on device, scalar special functions on host
    sigma, coef_c, su_c = ag(sigma_l), ag(coef_c_l), ag(su_c_l)
    df_res = float(S - k)
    s2 = sigma * sigma
    med = float(torch.median(s2))
    x = torch.clamp(s2, min=1e-5 * (med if med > 0 else 1.0))
    e = torch.log(x) - float(special.digamma(df_res / 2)) + np.log(df_res / 2)
    emean = float(e.mean())
    evar = float(((e - emean) ** 2).sum() / (G - 1)) - float(special.polygamma(1, df_res / 2))
    if evar > 0:
        df0 = 2 * float(O.trigamma_inverse(evar)[0])
        s20 = float(np.exp(emean + special.digamma(df0 / 2) - np.log(df0 / 2)))
    else:
        df0, s20 = float("inf"), float(np.exp(emean))
    df_cap = df_res * G
    s2_post = (df_res * s2 + df0 * s20) / (df_res + df0) if np.isfinite(df0) else torch.full_like(s2, s20)
    t_pre = coef_c / su_c / s2_post.sqrt()
    df_total = min(df_res + df0, df_cap)
    ntarget = int(np.ceil(proportion / 2 * G))
    top = torch.topk(t_pre.abs(), ntarget)
    v0_lim = np.array(stdev_coef_lim) ** 2 / s20
    # only the ntarget largest |t| enter limma's tmixture.vector; the gene count enters its rank formula
    var_prior = _tmixture_top(top.values.cpu().numpy(), su_c[top.indices].cpu().numpy(), df_total, proportion, v0_lim, G)
    if not np.isfinite(var_prior):
        var_prior = 1.0 / s20
    t_l = torch.empty(Gl, **f64)
    p_l = torch.empty(Gl, **f64)
    lods_l = torch.empty(Gl, **f64)
    s2p = torch.empty(Gl, **f64)
    dft = torch.empty(Gl, **f64)
    if Gl:
        _native.check(lb.cml_ebayes(P(coef_c_l), P(su_c_l), P(sigma_l), None, df_res, Gl, s20, df0, df_cap, var_prior, proportion,
                                    P(t_l), P(p_l), P(lods_l), P(s2p), P(dft), st), "cml_ebayes")
    t, p, lods = ag(t_l), ag(p_l), ag(lods_l)
    # ---- BH + topTable (device sort + cummin)
    ordp = torch.argsort(p, descending=True, stable=True)
    ranks = G - torch.arange(G, device=dev, dtype=torch.float64)
    adj_sorted = torch.cummin(p[ordp] * G / ranks, 0).values.clamp(max=1.0)
    adj = torch.empty_like(p)
    adj[ordp] = adj_sorted
    sig = (adj < p_value) & (coef_c.abs() > lfc)
    idx = torch.nonzero(sig).flatten()
    idx = idx[torch.argsort(p[idx], stable=True)]
    return {"coef_c": coef_c, "amean": amean, "t": t, "p": p, "adj": adj, "lods": lods, "s20": s20, "df0": df0, "var_prior": float(var_prior),
            "idx": idx, "trend": (sx, sy, ux, uy)}


def _tmixture_top(t_top: np.ndarray, su_top: np.ndarray, df: float, proportion: float, v0_lim, ngenes: int) -> float:
    """limma ``tmixture.vector`` given only the ``ntarget`` largest |t| (already sorted descending)."""
    from scipy import stats
    ntarget = t_top.size
    if ntarget < 1:
        return float("nan")
    p = max(ntarget / ngenes, proportion)
    v1 = su_top ** 2
    r = np.arange(1, ntarget + 1)
    p0 = 2 * stats.t.sf(t_top, df)
    ptarget = ((r - 0.5) / ngenes - (1 - p) * p0) / p
    v0 = np.zeros(ntarget)
    pos = ptarget > p0
    if pos.any():
        q = stats.t.isf(ptarget[pos] / 2, df)
        v0[pos] = v1[pos] * ((t_top[pos] / q) ** 2 - 1)
    v0 = np.minimum(np.maximum(v0, v0_lim[0]), v0_lim[1])
    return float(v0.mean())


def _oracle_general(c, design, contrast, min_frac):
    n = c.shape[1]
    keep = O.cpm_filter(c, 1.0, min_frac * n, strict=True)
    x = c[keep]
    lib = c.sum(0)
    nf = O.tmm_factors(x, lib)
    e, w, trend = O.voom(x, design, lib * nf)
    fit = O.lm_fit(e, design, w)
    coef, su = O.contrasts_fit(fit, contrast)
    eb = O.ebayes(coef, su, fit.sigma, fit.df_residual)
    idx, table = O.top_table(coef, fit.amean, eb)
    return {"keep": keep, "norm_factors": nf, "lib_size": lib, "E": e, "weights": w, "trend": trend, "fit": fit,
            "coef": coef, "stdev_unscaled": su, "ebayes": eb, "deg_index": idx, "table": table}


def write_deg_table(table: pd.DataFrame, path: str) -> None:
    """CSV in the reference's layout: ``"",logFC,AveExpr,t,P.Value,adj.P.Val,B`` with gene ids as row names
    (JSmith_code/Results/TARGET_AML_High.Std.Risk_vs_LowRisk_DEGs.csv, written at Differential_Expression_and_Lasso.Rmd:261)."""
    import csv
    with open(path, "w", newline="") as fh:
        w = csv.writer(fh, quoting=csv.QUOTE_NONNUMERIC)
        w.writerow([""] + list(table.columns))
        for gid, row in table.iterrows():
            w.writerow([gid] + [float(v) for v in row.values])
