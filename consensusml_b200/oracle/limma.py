"""Float64 NumPy/SciPy restatement of the edgeR + limma calls the reference makes.

Pipeline restated: ``DGEList -> cpm filter -> calcNormFactors (TMM) -> cpm(log=TRUE, prior.count=1)``
(JSmith_code/Normalization_and_Batch_Effect_Investigation.Rmd:49-63;
JSmith_code/Differential_Expression_and_Lasso.Rmd:217-227) and ``voom -> lmFit -> contrasts.fit -> eBayes
-> topTable`` (JSmith_code/Limma_Voom_DE_Function.R:25-42).  The reference contains only the *calls*; the
algorithms below follow the published descriptions of edgeR's TMM (Robinson & Oshlack 2010), limma's
voom (Law et al. 2014), Cleveland's lowess and Smyth's empirical-Bayes moderation (2004).  This module is the
oracle for the CUDA kernels in csrc/norm.cu, csrc/voom.cu, csrc/ebayes.cu and the timed CPU arm of the
normalisation / DE benchmarks.  Matrices are genes x samples, like the reference.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np
from scipy import special, stats


# --------------------------------------------------------------------------------------- CPM / filter
def cpm(counts: np.ndarray, lib_size: Optional[np.ndarray] = None) -> np.ndarray:
    lib = counts.sum(0) if lib_size is None else lib_size
    return counts / lib[None, :] * 1e6


def cpm_filter(counts: np.ndarray, min_cpm: float = 1.0, min_samples: float = 5, strict: bool = False) -> np.ndarray:
    """``rowSums(cpm(dge) >= 1) >= k`` (``strict`` -> ``> k``, the voom_DE variant, Limma_Voom_DE_Function.R:27)."""
    n_ok = (cpm(counts) >= min_cpm).sum(1)
    return n_ok > min_samples if strict else n_ok >= min_samples


# --------------------------------------------------------------------------------------- TMM
def _rank_avg(v: np.ndarray) -> np.ndarray:
    return stats.rankdata(v, method="average")


def _tmm_one(obs, ref, n_obs, n_ref, logratio_trim=0.3, sum_trim=0.05, a_cutoff=-1e10) -> float:
    with np.errstate(divide="ignore", invalid="ignore"):
        log_r = np.log2((obs / n_obs) / (ref / n_ref))
        abs_e = (np.log2(obs / n_obs) + np.log2(ref / n_ref)) / 2
        v = (n_obs - obs) / n_obs / obs + (n_ref - ref) / n_ref / ref
    fin = np.isfinite(log_r) & np.isfinite(abs_e) & (abs_e > a_cutoff)
    log_r, abs_e, v = log_r[fin], abs_e[fin], v[fin]
    if log_r.size == 0 or np.max(np.abs(log_r)) < 1e-6:
        return 1.0
    n = log_r.size
    lo_l = np.floor(n * logratio_trim) + 1
    hi_l = n + 1 - lo_l
    lo_s = np.floor(n * sum_trim) + 1
    hi_s = n + 1 - lo_s
    rl, rs = _rank_avg(log_r), _rank_avg(abs_e)
    keep = (rl >= lo_l) & (rl <= hi_l) & (rs >= lo_s) & (rs <= hi_s)
    f = np.sum(log_r[keep] / v[keep]) / np.sum(1 / v[keep])
    if not np.isfinite(f):
        f = 0.0
    return float(2.0 ** f)


def tmm_ref_column(counts: np.ndarray, lib_size: np.ndarray) -> int:
    y = counts / lib_size[None, :]
    f75 = np.quantile(y, 0.75, axis=0)          # R type-7 quantile == NumPy's default 'linear'
    if np.median(f75) < 1e-20:
        return int(np.argmax(np.sqrt(counts).sum(0)))
    return int(np.argmin(np.abs(f75 - f75.mean())))


def tmm_factors(counts: np.ndarray, lib_size: Optional[np.ndarray] = None, ref_column: Optional[int] = None,
                normalize: bool = True) -> np.ndarray:
    """edgeR ``calcNormFactors(method="TMM")``: trimmed mean of M-values against a reference sample,
    precision weighted, scaled to geometric mean 1."""
    counts = np.asarray(counts, dtype=np.float64)
    lib = counts.sum(0) if lib_size is None else np.asarray(lib_size, dtype=np.float64)
    x = counts[(counts > 0).any(1)]
    ref = tmm_ref_column(x, lib) if ref_column is None else ref_column
    f = np.array([_tmm_one(x[:, i], x[:, ref], lib[i], lib[ref]) for i in range(x.shape[1])])
    return f / np.exp(np.mean(np.log(f))) if normalize else f


def log_cpm(counts: np.ndarray, lib_size: np.ndarray, norm_factors: Optional[np.ndarray] = None,
            prior_count: float = 1.0, mean_lib: Optional[float] = None) -> np.ndarray:
    """edgeR ``cpm(dge, log=TRUE, prior.count=p)`` - the prior is scaled by relative library size (``mean_lib``: the
    mean effective library size when ``counts`` holds only a shard of the samples)."""
    lib = lib_size * (1.0 if norm_factors is None else norm_factors)
    pcs = lib / (lib.mean() if mean_lib is None else mean_lib) * prior_count
    return np.log2((counts + pcs[None, :]) / (lib + 2 * pcs)[None, :] * 1e6)


# --------------------------------------------------------------------------------------- lowess
def lowess(x: np.ndarray, y: np.ndarray, f: float = 2.0 / 3.0, iters: int = 3, delta: Optional[float] = None):
    """Cleveland's robust locally weighted regression as in R's ``lowess`` (x must be sorted ascending).
    Returns fitted values at every x."""
    x = np.asarray(x, float)
    y = np.asarray(y, float)
    n = x.size
    if delta is None:
        delta = 0.01 * (x[-1] - x[0])
    ns = int(max(min(int(f * n + 1e-7), n), 2))
    ys = np.zeros(n)
    rw = np.ones(n)
    res = np.zeros(n)
    xrange = x[-1] - x[0]
    for it in range(iters + 1):
        nleft, nright, last, i = 0, ns - 1, -1, 0
        while True:
            if nright < n - 1:
                d1 = x[i] - x[nleft]
                d2 = x[nright + 1] - x[i]
                if d1 > d2:
                    nleft += 1
                    nright += 1
                    continue
            xs = x[i]
            h = max(xs - x[nleft], x[nright] - xs)
            h9, h1 = 0.999 * h, 0.001 * h
            # weights over [nleft, nrt]
            j = nleft
            hi = nleft
            while hi < n and (abs(x[hi] - xs) <= h9 or x[hi] <= xs):
                hi += 1
            seg = slice(nleft, hi)
            r = np.abs(x[seg] - xs)
            w = np.where(r <= h9, np.where(r <= h1, 1.0, (1 - (r / h) ** 3) ** 3 if h > 0 else 1.0), 0.0)
            if it > 0:
                w = w * rw[seg]
            a = w.sum()
            if a <= 0:
                ys[i] = y[i]
            else:
                w = w / a
                if h > 0:
                    a1 = np.sum(w * x[seg])
                    b = xs - a1
                    c = np.sum(w * (x[seg] - a1) ** 2)
                    if np.sqrt(c) > 0.001 * xrange:
                        b /= c
                        w = w * (b * (x[seg] - a1) + 1)
                ys[i] = np.sum(w * y[seg])
            if last < i - 1:
                denom = x[i] - x[last]
                jj = np.arange(last + 1, i)
                alpha = (x[jj] - x[last]) / denom
                ys[jj] = alpha * ys[i] + (1 - alpha) * ys[last]
            last = i
            cut = x[last] + delta
            i = last + 1
            while i < n:
                if x[i] > cut:
                    break
                if x[i] == x[last]:
                    ys[i] = ys[last]
                    last = i
                i += 1
            i = max(last + 1, i - 1)
            if last >= n - 1:
                break
        res = y - ys
        sc = np.abs(res).sum() / n
        if it >= iters:
            break
        ar = np.abs(res)
        m1 = n // 2
        srt = np.sort(ar)
        if n % 2 == 0:
            cmad = 3.0 * (srt[m1] + srt[n - m1 - 1])
        else:
            cmad = 6.0 * srt[m1]
        if cmad < 1e-7 * sc:
            break
        c9, c1 = 0.999 * cmad, 0.001 * cmad
        rw = np.where(ar <= c1, 1.0, np.where(ar <= c9, (1 - (ar / cmad) ** 2) ** 2, 0.0))
    return ys


def lowess_eval_points(x: np.ndarray, delta: float) -> np.ndarray:
    """Indices at which lowess performs a local fit (the rest are interpolated / tied copies)."""
    n = x.size
    pts = []
    last, i = -1, 0
    while True:
        pts.append(i)
        last = i
        cut = x[last] + delta
        i = last + 1
        while i < n:
            if x[i] > cut:
                break
            if x[i] == x[last]:
                last = i
            i += 1
        i = max(last + 1, i - 1)
        if last >= n - 1:
            break
    return np.asarray(pts, dtype=np.int64)


# --------------------------------------------------------------------------------------- voom / lmFit
@dataclass
class LinearFit:
    coefficients: np.ndarray      # genes x k
    stdev_unscaled: np.ndarray    # genes x k
    sigma: np.ndarray             # genes
    df_residual: np.ndarray       # genes
    amean: np.ndarray             # genes
    xtwx_inv: Optional[np.ndarray] = None   # genes x k x k
    design: Optional[np.ndarray] = None


def lm_fit(y: np.ndarray, design: np.ndarray, weights: Optional[np.ndarray] = None) -> LinearFit:
    """Per-gene (weighted) least squares - limma ``lmFit`` with ``method="ls"``."""
    g, n = y.shape
    k = design.shape[1]
    w = np.ones_like(y) if weights is None else weights
    xtwx = np.einsum("ni,gn,nj->gij", design, w, design)
    xtwy = np.einsum("ni,gn->gi", design, w * y)
    inv = np.linalg.inv(xtwx)
    coef = np.einsum("gij,gj->gi", inv, xtwy)
    res = y - coef @ design.T
    df = np.full(g, n - k, dtype=np.float64)
    sigma = np.sqrt((w * res * res).sum(1) / df)
    su = np.sqrt(np.einsum("gii->gi", inv))
    return LinearFit(coef, su, sigma, df, y.mean(1), inv, design)


def voom(counts: np.ndarray, design: np.ndarray, lib_size: np.ndarray, span: float = 0.5):
    """limma ``voom``: log2-cpm with +0.5 / +1 offsets, mean-variance lowess trend, precision weights.
    Returns (E, weights, (sx, sy, trend_x, trend_y))."""
    y = np.log2((counts + 0.5) / (lib_size + 1.0)[None, :] * 1e6)
    fit = lm_fit(y, design)
    sx = fit.amean + np.mean(np.log2(lib_size + 1.0)) - np.log2(1e6)
    sy = np.sqrt(fit.sigma)
    allzero = counts.sum(1) == 0
    keep = ~allzero
    o = np.argsort(sx[keep], kind="stable")
    lx, ly_in = sx[keep][o], sy[keep][o]
    ly = lowess(lx, ly_in, f=span)
    # approxfun(rule=2, ties=list("ordered", mean))
    ux, inv = np.unique(lx, return_inverse=True)
    uy = np.bincount(inv, weights=ly) / np.bincount(inv)
    fitted = fit.coefficients @ design.T
    fitted_logcount = np.log2(1e-6 * (2.0 ** fitted) * (lib_size + 1.0)[None, :])
    tr = np.interp(fitted_logcount, ux, uy)       # constant extrapolation == rule=2
    w = 1.0 / tr ** 4
    return y, w, (sx, sy, ux, uy)


def contrasts_fit(fit: LinearFit, contrast: np.ndarray):
    """Apply a single contrast vector c: beta_c = c'beta; unscaled sd = sqrt(c'(X'WX)^-1 c).
    (For the group-means design the reference uses, limma's orthogonal shortcut gives the same number.)"""
    c = np.asarray(contrast, float)
    coef = fit.coefficients @ c
    su = np.sqrt(np.einsum("i,gij,j->g", c, fit.xtwx_inv, c))
    return coef, su


def trigamma_inverse(x: np.ndarray) -> np.ndarray:
    x = np.atleast_1d(np.asarray(x, float))
    y = 0.5 + 1.0 / x
    for _ in range(50):
        tri = special.polygamma(1, y)
        dif = tri * (1 - tri / x) / special.polygamma(2, y)
        y = y + dif
        if np.max(-dif / y) < 1e-8:
            break
    y = np.where(x > 1e7, 1 / np.sqrt(x), y)
    y = np.where(x < 1e-6, 1 / x, y)
    return y


def fit_f_dist(var: np.ndarray, df1: np.ndarray) -> Tuple[float, float]:
    """Moment estimation of the scaled-F prior on the variances (limma ``fitFDist``)."""
    ok = np.isfinite(var) & np.isfinite(df1) & (var > -1e-15) & (df1 > 1e-15)
    x, d = np.maximum(var[ok], 0), df1[ok]
    m = np.median(x)
    if m == 0:
        m = 1.0
    x = np.maximum(x, 1e-5 * m)
    e = np.log(x) - special.digamma(d / 2) + np.log(d / 2)
    emean = e.mean()
    evar = np.sum((e - emean) ** 2) / (e.size - 1) - np.mean(special.polygamma(1, d / 2))
    if evar > 0:
        df2 = 2 * float(trigamma_inverse(evar)[0])
        s20 = float(np.exp(emean + special.digamma(df2 / 2) - np.log(df2 / 2)))
    else:
        df2 = np.inf
        s20 = float(np.exp(emean))
    return s20, df2


def tmixture_vector(tstat, stdev_unscaled, df, proportion, v0_lim=None) -> float:
    ngenes = tstat.size
    ntarget = int(np.ceil(proportion / 2 * ngenes))
    if ntarget < 1:
        return np.nan
    p = max(ntarget / ngenes, proportion)
    tstat = np.abs(tstat)
    max_df = np.max(df)
    i = df < max_df
    if i.any():
        tail = stats.t.logsf(tstat[i], df[i])
        tstat = tstat.copy()
        tstat[i] = stats.t.isf(np.exp(tail), max_df)
    o = np.argsort(-tstat, kind="stable")[:ntarget]
    ts, v1 = tstat[o], stdev_unscaled[o] ** 2
    r = np.arange(1, ntarget + 1)
    p0 = 2 * stats.t.sf(ts, max_df)
    ptarget = ((r - 0.5) / ngenes - (1 - p) * p0) / p
    v0 = np.zeros(ntarget)
    pos = ptarget > p0
    if pos.any():
        qtarget = stats.t.isf(ptarget[pos] / 2, max_df)
        v0[pos] = v1[pos] * ((ts[pos] / qtarget) ** 2 - 1)
    if v0_lim is not None:
        v0 = np.minimum(np.maximum(v0, v0_lim[0]), v0_lim[1])
    return float(v0.mean())


@dataclass
class EBayesResult:
    t: np.ndarray
    p_value: np.ndarray
    lods: np.ndarray
    s2_post: np.ndarray
    df_total: np.ndarray
    s2_prior: float
    df_prior: float
    var_prior: float


def ebayes(coef, stdev_unscaled, sigma, df_residual, proportion=0.01, stdev_coef_lim=(0.1, 4.0)) -> EBayesResult:
    """limma ``eBayes``: squeeze variances, moderated t, two-sided p, B-statistic (log-odds)."""
    s2 = sigma ** 2
    s20, df0 = fit_f_dist(s2, df_residual)
    if np.isfinite(df0):
        s2_post = (df_residual * s2 + df0 * s20) / (df_residual + df0)
    else:
        s2_post = np.full_like(s2, s20)
    t = coef / stdev_unscaled / np.sqrt(s2_post)
    df_total = np.minimum(df_residual + df0, df_residual.sum())
    p = 2 * stats.t.sf(np.abs(t), df_total)
    v0_lim = np.array(stdev_coef_lim) ** 2 / s20
    var_prior = tmixture_vector(t, stdev_unscaled, df_total, proportion, v0_lim)
    if not np.isfinite(var_prior):
        var_prior = 1.0 / s20
    r = (stdev_unscaled ** 2 + var_prior) / stdev_unscaled ** 2
    t2 = t * t
    if np.isfinite(df0) and df0 <= 1e6:
        kernel = (1 + df_total) / 2 * np.log((t2 + df_total) / (t2 / r + df_total))
    else:
        kernel = t2 * (1 - 1 / r) / 2
    lods = np.log(proportion / (1 - proportion)) - np.log(r) / 2 + kernel
    return EBayesResult(t, p, lods, s2_post, df_total, s20, df0, var_prior)


def bh_adjust(p: np.ndarray) -> np.ndarray:
    """Benjamini-Hochberg adjusted p-values (R ``p.adjust(method="BH")``)."""
    n = p.size
    o = np.argsort(-p, kind="stable")
    ranks = n - np.arange(n)
    adj = np.minimum.accumulate(p[o] * n / ranks)
    out = np.empty(n)
    out[o] = np.minimum(adj, 1.0)
    return out


def top_table(coef, amean, eb: EBayesResult, p_value=0.05, lfc=1.0):
    """limma ``topTable(p.value=, lfc=, adjust.method="BH", sort.by="P")``: returns (row indices, table dict)."""
    adj = bh_adjust(eb.p_value)
    sig = (adj < p_value) & (np.abs(coef) > lfc)
    idx = np.flatnonzero(sig)
    idx = idx[np.argsort(eb.p_value[idx], kind="stable")]
    return idx, {"logFC": coef[idx], "AveExpr": amean[idx], "t": eb.t[idx], "P.Value": eb.p_value[idx],
                 "adj.P.Val": adj[idx], "B": eb.lods[idx]}


def voom_de(counts: np.ndarray, group: np.ndarray, min_frac: float = 0.05):
    """The reference's ``voom_DE``: ``group`` is 1 for the non-reference level, 0 for the reference.
    Returns dict with kept-gene mask, norm factors, fit pieces and the DEG table indices (into kept genes)."""
    counts = np.asarray(counts, float)
    n = counts.shape[1]
    keep = cpm_filter(counts, 1.0, min_frac * n, strict=True)
    x = counts[keep]
    lib = counts.sum(0)                       # DGEList subsetting keeps the original library sizes
    nf = tmm_factors(x, lib)
    design = np.stack([group == 1, group == 0], 1).astype(float)   # ~0+group, non-reference level first
    e, w, trend = voom(x, design, lib * nf)
    fit = lm_fit(e, design, w)
    coef, su = contrasts_fit(fit, np.array([1.0, -1.0]))
    eb = ebayes(coef, su, fit.sigma, fit.df_residual)
    idx, table = top_table(coef, fit.amean, eb)
    return {"keep": keep, "norm_factors": nf, "lib_size": lib, "E": e, "weights": w, "trend": trend, "fit": fit,
            "coef": coef, "stdev_unscaled": su, "ebayes": eb, "deg_index": idx, "table": table}
