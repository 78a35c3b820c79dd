"""DESeq2's dispersion estimation, variance-stabilising transformation and regularised log, as MLSeq uses them for its
"deseq-vst" / "deseq-rlog" pre-processing (VikasP/AML.R:174-264: DESeqDataSetFromMatrix, classify(..., preProcessing = ...)).

What is restated here, step for step (DESeq2 1.2x; the reference calls it through MLSeq, so no R source of it is in the repo):

  estimateDispersionsGeneEst   linear-model means of a group-indicator design (``~ 1`` when blind), start value
                               min(rough, moments) estimate, line search of the Cox-Reid adjusted profile likelihood over
                               log(alpha) with the Armijo rule of ``fitDisp``, the "no increase -> keep the start" and
                               "unconverged -> two-level grid" rules, clamp to [1e-8, max(10, n)]
  estimateDispersionsFit       ``parametric``: alpha(mu) = asymptDisp + extraPois / mu by an iterated Gamma-family GLM with
                               identity link on the genes with residual ratio in (1e-4, 15); ``mean``: trimmed mean;
                               ``local``: a lowess of log(alpha) on log(mu) stands in for locfit
  varianceStabilizingTransformation   closed forms for the parametric and the mean trend; numerical integration of
                               1 / sqrt(mu + alpha(mu) mu^2) on an asinh grid for the local trend
  rlog                         prior variance by matching the weighted upper 5 % quantile of the log fold changes to a
                               normal, ridge NB GLM with one coefficient per sample (``fitBeta``: IRLS, deviance
                               criterion 1e-4, at most 100 iterations); "frozen" application to new samples with the
                               training intercept and prior variance

The per-gene fits are kernels (csrc/deseq.cu: one warp per gene); on CPU tensors outside ``_native.emulation()`` the same
algorithms run vectorised over genes in float64 torch, which is also the oracle of the kernel tests."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np
import torch

from .. import _native

MIN_DISP = 1e-8
_native.register_signature("cml_nb_dispersion", [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                                   ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                                   ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                   ctypes.c_void_p])
_native.register_signature("cml_rlog", [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                          ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                          ctypes.c_void_p, ctypes.c_void_p])


@dataclass
class DispersionFit:
    size_factors: torch.Tensor       # [S]
    loggeo: torch.Tensor             # [G] mean log count of the genes without zeros (NaN otherwise): freezes the size factors
    base_mean: torch.Tensor          # [G]
    base_var: torch.Tensor           # [G]
    disp_gene: torch.Tensor          # [G] gene-wise estimates (NaN: all-zero gene)
    disp_fit: torch.Tensor           # [G] trend at the gene's mean
    fit_type: str
    coefs: Tuple[float, float]       # (asymptDisp, extraPois) for "parametric", (mean, 0) for "mean"
    iters: torch.Tensor              # [G] line-search iterations
    trend: Optional[tuple] = None    # "local": (log mean grid, log dispersion on it)
    vst_table: Optional[tuple] = None  # "local": the integrated transform tabulated on the data it was fitted to (asinh grid, values, scale, shift)

    def dispersion_at(self, mu: torch.Tensor) -> torch.Tensor:
        if self.fit_type == "parametric":
            return self.coefs[0] + self.coefs[1] / mu
        if self.fit_type == "mean":
            return torch.full_like(mu, self.coefs[0])
        gx, gy = self.trend
        return torch.exp(_interp(torch.log(mu.clamp_min(1e-300)), gx, gy))


def _interp(x: torch.Tensor, gx: torch.Tensor, gy: torch.Tensor) -> torch.Tensor:
    """Piecewise-linear interpolation with constant extrapolation (gx ascending)."""
    idx = torch.searchsorted(gx, x.contiguous()).clamp(1, gx.numel() - 1)
    x0, x1, y0, y1 = gx[idx - 1], gx[idx], gy[idx - 1], gy[idx]
    t = ((x - x0) / (x1 - x0)).clamp(0, 1)
    return y0 + t * (y1 - y0)


# ------------------------------------------------------------------------------------------------ size factors
def size_factors(counts: torch.Tensor, loggeo: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Median-of-ratios size factors and the per-gene mean log counts they are taken against.  With ``loggeo`` given (the
    training set's) the new samples are normalised against it - ``estimateSizeFactors(geoMeans = )``."""
    c = counts.double()
    logc = torch.log(c)
    if loggeo is None:
        ok = (c > 0).all(1)
        loggeo = torch.where(ok, logc.mean(1), torch.full((c.shape[0],), float("nan"), dtype=torch.float64, device=c.device))
    sf = torch.empty(c.shape[1], dtype=torch.float64, device=c.device)
    for j in range(c.shape[1]):                   # per-sample filter: finite reference AND a positive count in this sample
        use = torch.isfinite(loggeo) & (c[:, j] > 0)
        sf[j] = torch.exp(torch.quantile((logc[:, j] - loggeo)[use], 0.5)) if bool(use.any()) else 1.0
    return sf, loggeo


# ------------------------------------------------------------------------------------------------ gene-wise dispersions
def _apl(log_alpha, y, mu, onehot, want_grad):
    """Cox-Reid adjusted profile log-likelihood of log(alpha) for every gene ([G]) and its derivative."""
    alpha = torch.exp(log_alpha)[:, None]
    an1 = 1.0 / alpha
    ll = (torch.lgamma(y + an1) - torch.lgamma(an1) - y * torch.log(mu + an1) - an1 * torch.log1p(mu * alpha)).sum(1)
    w = 1.0 / (1.0 / mu + alpha)
    B = w @ onehot                                                # [G, n_groups]
    val = ll - 0.5 * torch.log(B).sum(1)
    if not want_grad:
        return val, None
    dll = (torch.digamma(an1) + torch.log1p(mu * alpha) - mu * alpha / (1.0 + mu * alpha) - torch.digamma(y + an1) + y / (mu + an1)).sum(1)
    dB = -(w * w) @ onehot
    grad = (dll * an1[:, 0] ** 2 - 0.5 * (dB / B).sum(1)) * alpha[:, 0]
    return val, grad


def _gene_dispersions_torch(c: torch.Tensor, sf: torch.Tensor, group: torch.Tensor, n_groups: int, min_disp: float, max_disp: float):
    G, S = c.shape
    q = c / sf[None, :]
    bm = q.mean(1)
    bv = q.var(1, unbiased=True)
    onehot = torch.nn.functional.one_hot(group.long(), n_groups).double()
    gmean = (q @ onehot) / onehot.sum(0)[None, :]
    mu_norm = gmean @ onehot.t()                                  # linear-model fit of the normalised counts
    mr = mu_norm.clamp_min(1.0)
    rough = ((((q - mr) ** 2 - mr) / mr ** 2).sum(1) / (S - n_groups)).clamp_min(0)
    moments = (bv - (1.0 / sf).mean() * bm) / (bm * bm)
    alive = bm > 0
    alpha0 = torch.minimum(rough, moments).clamp(min_disp, max_disp)
    alpha0 = torch.where(alive, alpha0, torch.ones_like(alpha0))
    mu = (mu_norm * sf[None, :]).clamp_min(0.5)
    y = c
    eps, kappa0, tol, maxit = 1e-4, 1.0, 1e-6, 100
    min_log_alpha = float(np.log(min_disp / 10.0))
    a = torch.log(alpha0)
    lp, dlp = _apl(a, y, mu, onehot, True)
    initial_lp = lp.clone()
    kappa = torch.full_like(a, kappa0)
    iters = torch.zeros(G, dtype=torch.int64)
    accepts = torch.zeros(G, dtype=torch.int64)
    active = alive.clone()
    for _ in range(maxit):
        if not bool(active.any()):
            break
        iters += active.long()
        prop = a + kappa * dlp
        kappa = torch.where(active & (prop < -30.0), (-30.0 - a) / dlp, kappa)
        prop = a + kappa * dlp
        kappa = torch.where(active & (prop > 10.0), (10.0 - a) / dlp, kappa)
        prop = a + kappa * dlp
        theta_kappa = -_apl(torch.where(active, prop, a), y, mu, onehot, False)[0]
        theta_hat = -lp - kappa * eps * dlp * dlp
        acc = active & (theta_kappa <= theta_hat)
        rej = active & ~acc
        accepts += acc.long()
        a = torch.where(acc, prop, a)
        lpn, dn = _apl(a, y, mu, onehot, True)
        change = lpn - lp
        lp = torch.where(acc, lpn, lp)
        stop = acc & ((change < tol) | (a < min_log_alpha))
        go = acc & ~stop
        dlp = torch.where(go, dn, dlp)
        kappa = torch.where(go, torch.minimum(kappa * 1.1, torch.full_like(kappa, kappa0)), kappa)
        kappa = torch.where(go & (accepts % 5 == 0), kappa * 0.5, kappa)
        kappa = torch.where(rej, kappa * 0.5, kappa)
        active = active & ~stop
    disp = torch.exp(a)
    disp = torch.where(lp < initial_lp + initial_lp.abs() / 1e6, alpha0, disp)
    conv = (iters < maxit) & (iters != 1)
    refit = alive & ~conv & (disp > min_disp * 10.0)
    if bool(refit.any()):
        idx = torch.nonzero(refit).flatten()
        lo, hi = float(np.log(min_disp)), float(np.log(max_disp))
        grid = torch.linspace(lo, hi, 20, dtype=torch.float64)
        vals = torch.stack([_apl(torch.full((idx.numel(),), float(g), dtype=torch.float64), y[idx], mu[idx], onehot, False)[0] for g in grid], 1)
        abest = grid[vals.argmax(1)]
        d1 = float(grid[1] - grid[0])
        fine = torch.stack([abest - d1 + 2.0 * d1 / 19.0 * t for t in range(20)], 1)
        vals2 = torch.stack([_apl(fine[:, t], y[idx], mu[idx], onehot, False)[0] for t in range(20)], 1)
        disp[idx] = torch.exp(fine[torch.arange(idx.numel()), vals2.argmax(1)])
    disp = disp.clamp(min_disp, max_disp)
    disp = torch.where(alive, disp, torch.full_like(disp, float("nan")))
    return bm, bv, disp, iters.to(torch.int32)


def gene_dispersions(counts: torch.Tensor, sf: torch.Tensor, groups: Optional[torch.Tensor] = None):
    """(base_mean, base_var, disp_gene, iters) - DESeq2 ``estimateDispersionsGeneEst`` for the design ``~ groups`` (``~ 1`` if None)."""
    c = counts.double().contiguous()
    G, S = c.shape
    if groups is None:
        group = torch.zeros(S, dtype=torch.int32, device=c.device)
        n_groups = 1
    else:
        _, inv = torch.unique(groups.to(c.device), return_inverse=True)
        group = inv.to(torch.int32)
        n_groups = int(group.max()) + 1
    max_disp = float(max(10, S))
    if _native.on_device(c) and n_groups <= 16:
        sf = sf.double().contiguous().to(c.device)
        bm, bv, disp = (torch.empty(G, dtype=torch.float64, device=c.device) for _ in range(3))
        iters = torch.empty(G, dtype=torch.int32, device=c.device)
        P = _native.ptr
        rc = _native.cuda().cml_nb_dispersion(P(c), G, S, P(sf), P(group.contiguous()), n_groups, MIN_DISP, max_disp, P(bm), P(bv), P(disp),
                                              P(iters), _native.stream_ptr())
        _native.check(rc, "cml_nb_dispersion")
        return bm, bv, disp, iters
    return _gene_dispersions_torch(c, sf.double(), group, n_groups, MIN_DISP, max_disp)


# ------------------------------------------------------------------------------------------------ dispersion trend
def parametric_trend(means: torch.Tensor, disps: torch.Tensor) -> Tuple[float, float]:
    """``parametricDispersionFit``: Gamma GLM (identity link) of dispersion on 1 / mean, refitted on the genes whose ratio to
    the current fit lies in (1e-4, 15) until the coefficients settle.  Returns (asymptDisp, extraPois); raises if it does not
    converge or leaves the positive quadrant (DESeq2 then falls back to the local fit)."""
    m, d = means.double(), disps.double()
    coefs = torch.tensor([0.1, 1.0], dtype=torch.float64)
    for it in range(11):
        resid = d / (coefs[0] + coefs[1] / m)
        good = (resid > 1e-4) & (resid < 15)
        x = torch.stack([torch.ones_like(m[good]), 1.0 / m[good]], 1)
        yv = d[good]
        b = coefs.clone()
        for _ in range(25):                                       # IRLS of glm(family = Gamma(link = "identity"), start = coefs)
            fit = x @ b
            if bool((fit <= 0).any()):
                raise RuntimeError("parametric dispersion fit left the positive range")
            wt = 1.0 / (fit * fit)
            bn = torch.linalg.solve((x * wt[:, None]).t() @ x, (x * wt[:, None]).t() @ yv)
            done = float((bn - b).abs().max()) < 1e-8 * (float(b.abs().max()) + 1e-8)
            b = bn
            if done:
                break
        old, coefs = coefs, b
        if bool((coefs <= 0).any()):
            raise RuntimeError("parametric dispersion fit failed")
        if float((torch.log(coefs / old) ** 2).sum()) < 1e-6:
            return float(coefs[0]), float(coefs[1])
    raise RuntimeError("dispersion fit did not converge")


def _local_trend(means: torch.Tensor, disps: torch.Tensor):
    """log(alpha) against log(mean), smoothed (lowess, 2/3 span - in place of DESeq2's locfit) and tabulated on 200 points."""
    from ..oracle.limma import lowess as _lowess
    lx, ly = torch.log(means.double()).numpy(), torch.log(disps.double()).numpy()
    order = np.argsort(lx)
    sm = _lowess(lx[order], ly[order], f=2.0 / 3.0)
    gx = np.linspace(lx.min(), lx.max(), 200)
    return torch.from_numpy(gx), torch.from_numpy(np.interp(gx, lx[order], sm))


def estimate_dispersions(counts: torch.Tensor, sf: Optional[torch.Tensor] = None, groups: Optional[torch.Tensor] = None,
                         fit_type: str = "parametric") -> DispersionFit:
    """``estimateSizeFactors`` + ``estimateDispersionsGeneEst`` + ``estimateDispersionsFit``.  ``groups=None`` is the blind
    design ``~ 1`` that ``vst()`` / ``rlog()`` use by default."""
    loggeo = None
    if sf is None:
        sf, loggeo = size_factors(counts)
    else:
        loggeo = size_factors(counts)[1]
    bm, bv, dg, iters = gene_dispersions(counts, sf, groups)
    bm_c, dg_c = bm.cpu(), dg.cpu()
    use = torch.isfinite(dg_c) & (dg_c >= 100 * MIN_DISP)         # useForFit: estimates that did not collapse to the floor
    trend = None
    coefs = (float("nan"), 0.0)
    if fit_type == "parametric":
        try:
            coefs = parametric_trend(bm_c[use], dg_c[use])
        except RuntimeError:
            fit_type = "local"                                    # DESeq2's fallback
    if fit_type == "local":
        if int(use.sum()) >= 10:
            trend = tuple(t.to(counts.device) for t in _local_trend(bm_c[use], dg_c[use]))
        else:
            fit_type = "mean"
    if fit_type == "mean":
        ok = torch.isfinite(dg_c) & (dg_c > 10 * MIN_DISP)
        v = torch.sort(dg_c[ok]).values
        k = int(np.floor(v.numel() * 0.001))                      # mean(..., trim = 0.001)
        coefs = (float(v[k: v.numel() - k].mean()) if v.numel() else 0.01, 0.0)
    out = DispersionFit(sf.to(counts.device), loggeo.to(counts.device), bm, bv, dg, bm, fit_type, coefs, iters, trend)
    out.disp_fit = torch.where(bm > 0, out.dispersion_at(bm.clamp_min(1e-300)), torch.full_like(bm, float("nan")))
    return out


# ------------------------------------------------------------------------------------------------ VST
def vst(counts: torch.Tensor, fit: Optional[DispersionFit] = None, *, sf: Optional[torch.Tensor] = None, groups: Optional[torch.Tensor] = None,
        fit_type: str = "parametric") -> Tuple[torch.Tensor, DispersionFit]:
    """``varianceStabilizingTransformation`` (log2-like scale).  With ``fit`` given (a training set's) the counts are
    normalised against its geometric means and transformed with ITS dispersion function - what MLSeq does to test samples."""
    if fit is None:
        fit = estimate_dispersions(counts, sf, groups, fit_type)
        sf = fit.size_factors
    else:
        sf = size_factors(counts, fit.loggeo)[0]
    q = counts.double() / sf[None, :]
    ln2 = float(np.log(2.0))
    if fit.fit_type == "parametric":
        a0, a1 = fit.coefs                                        # asymptDisp, extraPois
        out = torch.log((1 + a1 + 2 * a0 * q + 2 * torch.sqrt(a0 * q * (1 + a1 + a0 * q))) / (4 * a0)) / ln2
    elif fit.fit_type == "mean":
        al = fit.coefs[0]
        out = (2 * torch.asinh(torch.sqrt(al * q)) - np.log(al) - np.log(4.0)) / ln2
    else:
        # numerical integration of 1 / sqrt(mu + alpha(mu) mu^2) on a grid that is uniform in asinh, then rescaled so that
        # the transform agrees with log2 for large counts (DESeq2 getVarianceStabilizedData, local fit).  The table and its
        # scaling are made once, from the data the trend was fitted to, and reused for new samples.
        if fit.vst_table is None:
            qmax = float(q.max()) * 8.0
            xg = torch.sinh(torch.linspace(0.0, float(np.arcsinh(max(qmax, 1.0))), 2000, dtype=torch.float64))[1:].to(q.device)
            integrand = 1.0 / torch.sqrt(xg + fit.dispersion_at(xg) * xg * xg)
            splf_x = torch.asinh((xg[1:] + xg[:-1]) / 2)
            splf_y = torch.cumsum((xg[1:] - xg[:-1]) * (integrand[1:] + integrand[:-1]) / 2, 0)
            h1, h2 = torch.quantile(q.mean(1), torch.tensor([0.95, 0.999], dtype=torch.float64, device=q.device))
            e1, e2 = _interp(torch.asinh(h1)[None], splf_x, splf_y)[0], _interp(torch.asinh(h2)[None], splf_x, splf_y)[0]
            eta = (torch.log2(h2) - torch.log2(h1)) / (e2 - e1)
            fit.vst_table = (splf_x, splf_y, float(eta), float(torch.log2(h1) - eta * e1))
        splf_x, splf_y, eta, xi = fit.vst_table
        aq = torch.asinh(q).flatten()
        out = eta * _interp(aq, splf_x, splf_y) + xi
        beyond = aq > splf_x[-1]                                    # past the table: the transform runs parallel to log2 there
        out = torch.where(beyond, eta * splf_y[-1] + xi + (torch.log2(torch.sinh(aq).clamp_min(1e-300)) - torch.log2(torch.sinh(splf_x[-1]))), out)
        out = out.reshape(q.shape)
    return out, fit


# ------------------------------------------------------------------------------------------------ rlog
def match_weighted_upper_quantile_for_variance(x: torch.Tensor, weights: torch.Tensor, upper_quantile: float = 0.05) -> float:
    """Variance of the zero-centred normal whose upper ``upper_quantile`` point equals the weighted quantile of |x|
    (Hmisc ``wtd.quantile(normwt = TRUE)`` semantics: weights rescaled to sum to n, linear interpolation of order statistics)."""
    from scipy.stats import norm
    ax, w = x.abs().double().flatten().cpu(), weights.double().flatten().cpu()
    order = torch.argsort(ax)
    ax, w = ax[order], w[order]
    n = ax.numel()
    cw = torch.cumsum(w * (n / float(w.sum())), 0)
    pos = 1.0 + (n - 1) * (1.0 - upper_quantile)
    lo = max(float(np.floor(pos)), 1.0)
    hi = min(lo + 1.0, float(n))
    frac = pos % 1.0
    at = lambda k: float(ax[min(int(torch.searchsorted(cw, torch.tensor(k - 1e-9, dtype=torch.float64))), n - 1)])
    q = (1.0 - frac) * at(lo) + frac * at(hi)
    return float((q / norm.ppf(1.0 - upper_quantile / 2.0)) ** 2)


def _rlog_torch(c, sf, alpha, bm, lam, lam0, intercept, tol, maxit):
    """Ridge IRLS of ``fitBeta`` for the design [1 | I], vectorised over genes (closed-form normal equations)."""
    G, S = c.shape
    fixed = intercept is not None
    alive = torch.isfinite(intercept) if fixed else bm > 0
    a = torch.where(alive, alpha, torch.ones_like(alpha))[:, None]
    size = 1.0 / a
    b0 = (torch.where(alive, intercept, torch.zeros_like(bm)) if fixed else torch.log(torch.where(alive, bm, torch.ones_like(bm))))[:, None]
    eta = b0.expand(G, S).clone()
    cst = (torch.lgamma(c + size) - torch.lgamma(size) - torch.lgamma(c + 1.0)).sum(1)
    dev_old = torch.zeros(G, dtype=torch.float64)
    iters = torch.zeros(G, dtype=torch.int32)
    active = alive.clone()
    for t in range(maxit):
        if not bool(active.any()):
            break
        mu = (sf[None, :] * torch.exp(eta)).clamp_min(0.5)
        w = mu / (1.0 + a * mu)
        z = torch.log(mu / sf[None, :]) + (c - mu) / mu
        if not fixed:
            cc = w * lam / (w + lam)
            b0n = ((cc * z).sum(1) / (lam0 + cc.sum(1)))[:, None]
        else:
            b0n = b0
        bj = w * (z - b0n) / (w + lam)
        eta_n = b0n + bj
        m2 = (sf[None, :] * torch.exp(eta_n)).clamp_min(0.5)
        dev = -2.0 * ((size * torch.log(size / (size + m2)) + c * torch.log(m2 / (size + m2))).sum(1) + cst)
        big = (bj.abs() > 30).any(1) | (b0n[:, 0].abs() > 30)
        conv = (dev - dev_old).abs() / (dev.abs() + 0.1)
        upd = active[:, None]
        eta = torch.where(upd, eta_n, eta)
        b0 = torch.where(upd, b0n, b0)
        bad = active & (big | torch.isnan(conv))
        done = active & ~bad & (conv < tol) & (t > 0)
        iters = torch.where(bad, torch.full_like(iters, maxit), torch.where(done, torch.full_like(iters, t), iters))
        dev_old = torch.where(active, dev, dev_old)
        active = active & ~bad & ~done
    iters = torch.where(active, torch.full_like(iters, maxit), iters)
    ln2 = float(np.log(2.0))
    out = torch.where(alive[:, None], eta / ln2, torch.zeros_like(eta))
    return out, torch.where(alive, b0[:, 0] / ln2, torch.full_like(bm, float("nan"))), iters


@dataclass
class RlogFit:
    values: torch.Tensor             # [G, S] log2 scale
    intercept: torch.Tensor          # [G] log2 (NaN for all-zero genes): ``mcols(dds)$rlogIntercept``
    beta_prior_var: float
    dispersion: DispersionFit
    iters: torch.Tensor


def rlog(counts: torch.Tensor, fit: Optional[DispersionFit] = None, *, frozen: Optional[RlogFit] = None, groups: Optional[torch.Tensor] = None,
         fit_type: str = "parametric", beta_prior_var: Optional[float] = None) -> RlogFit:
    """``rlog()``.  ``frozen`` = the RlogFit of a training set: new samples are normalised against its geometric means and fitted
    with its intercepts, dispersion trend and prior variance (``rlog(intercept = , betaPriorVar = )``, MLSeq's treatment of test data)."""
    c = counts.double().contiguous()
    G, S = c.shape
    ln2 = float(np.log(2.0))
    if frozen is not None:
        fit = frozen.dispersion
        sf = size_factors(c, fit.loggeo)[0]
        bm = fit.base_mean
        beta_prior_var = frozen.beta_prior_var
        intercept = (frozen.intercept * ln2).contiguous()
    else:
        if fit is None:
            fit = estimate_dispersions(c, None, groups, fit_type)
        sf = fit.size_factors
        bm = fit.base_mean
        intercept = None
        if beta_prior_var is None:
            nz = bm > 0
            q = c[nz] / sf[None, :]
            lfc = torch.log2(q + 0.5) - torch.log2(bm[nz] + 0.5)[:, None]
            wts = 1.0 / (1.0 / bm[nz] + fit.disp_fit[nz])
            beta_prior_var = match_weighted_upper_quantile_for_variance(lfc.t(), wts[None, :].expand(S, -1))
    alpha = torch.where(torch.isfinite(fit.disp_fit), fit.disp_fit, torch.ones_like(fit.disp_fit)).contiguous()
    lam = (1.0 / beta_prior_var) / ln2 ** 2                       # lambda on the natural-log scale
    lam0 = 1e-6 / ln2 ** 2
    if _native.on_device(c):
        out = torch.empty(G, S, dtype=torch.float64, device=c.device)
        b0 = torch.empty(G, dtype=torch.float64, device=c.device)
        iters = torch.empty(G, dtype=torch.int32, device=c.device)
        P = _native.ptr
        rc = _native.cuda().cml_rlog(P(c), G, S, P(sf.double().contiguous()), P(alpha), P(bm.contiguous()), P(intercept) if intercept is not None else None,
                                     lam, lam0, 1e-4, 100, P(out), P(b0), P(iters), _native.stream_ptr())
        _native.check(rc, "cml_rlog")
    else:
        out, b0, iters = _rlog_torch(c, sf.double(), alpha, bm, lam, lam0, intercept, 1e-4, 100)
    return RlogFit(out, b0 if frozen is None else frozen.intercept, float(beta_prior_var), fit, iters)
