"""Count-based discriminant classifiers - the MLSeq / DESeq2 zoo of VikasP/AML.R.

Reference call sites: ``DESeqDataSetFromMatrix`` + ``classify(method = "svmRadial" | "pam" | "PLDA" | "PLDA2" |
"NBLDA" | "voomDLDA" | "voomNSC", preProcessing = "deseq-vst" | "deseq-rlog", control = trainControl("repeatedcv",
number = 5, repeats = 10))`` and ``selectedGenes`` (VikasP/AML.R:167-291), on raw counts of the top-100-variance
genes with a +1 pseudo-count (:157-171).

Everything is expressed as batched tensor reductions over genes x samples (the per-class sums are one matmul with
the class-indicator matrix); on the kernel path the two reductions that are not a GEMM - the median-of-ratios size
factors and the Poisson-LDA scores - are ``size_factors_kernel`` / ``plda_scores_kernel`` (csrc/discriminant.cu).
Gene-sharded multi-GPU runs only need the per-sample scores all-reduced.  Implemented from the published methods: DESeq median-of-ratios (Anders & Huber 2010),
Poisson LDA (Witten 2011), NBLDA (Dong et al. 2016), nearest shrunken centroids (Tibshirani et al. 2002), voom-based
NSC / DLDA (Zararsiz et al. 2017)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import ctypes

import numpy as np
import torch

from .. import _native

_I, _P = ctypes.c_int, ctypes.c_void_p
_native.register_signature("cml_deseq_size_factors", [_P, _I, _I, _P, _P, _P])
_native.register_signature("cml_plda_scores", [_P, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P])


# ------------------------------------------------------------------------------------------------ normalisation
def deseq_size_factors(counts: torch.Tensor) -> torch.Tensor:
    """Median-of-ratios size factors; ``counts`` is genes x samples.  Genes with a zero in any sample are skipped."""
    c = counts.double()
    if _native.on_device(c) and c.shape[0] <= 25600:
        c = c.contiguous()
        G, S = c.shape
        scratch = torch.empty(G, dtype=torch.float64, device=c.device)
        sf = torch.empty(S, dtype=torch.float64, device=c.device)
        rc = _native.cuda().cml_deseq_size_factors(_native.ptr(c), G, S, _native.ptr(scratch), _native.ptr(sf), _native.stream_ptr())
        _native.check(rc, "cml_deseq_size_factors")
        return sf
    ok = (c > 0).all(1)
    logc = torch.log(c[ok])
    loggeo = logc.mean(1, keepdim=True)
    return torch.exp(torch.quantile(logc - loggeo, 0.5, dim=0))      # R's median: mean of the two middle values


def dispersion_moments(counts: torch.Tensor, sf: torch.Tensor) -> torch.Tensor:
    """Per-gene method-of-moments NB dispersion: (var - mean * mean(1/s)) / mean^2, floored at 0."""
    q = counts.double() / sf[None, :]
    mu = q.mean(1)
    var = q.var(1, unbiased=True)
    return ((var - mu * (1.0 / sf).mean()) / (mu * mu).clamp_min(1e-12)).clamp_min(0)


def deseq_vst(counts: torch.Tensor, sf: Optional[torch.Tensor] = None, *, fit=None, fit_type: str = "parametric", return_fit: bool = False):
    """DESeq2's ``varianceStabilizingTransformation`` (blind design): gene-wise dispersions by the Cox-Reid adjusted profile
    likelihood, parametric trend by the Gamma GLM, closed-form transform (``models.deseq``; kernels in csrc/deseq.cu).
    ``fit`` = a training set's ``DispersionFit``: transform new samples with its geometric means and dispersion function."""
    from . import deseq as DQ
    out, fit = DQ.vst(counts, fit, sf=sf, fit_type=fit_type)
    return (out, fit) if return_fit else out


def deseq_rlog(counts: torch.Tensor, sf: Optional[torch.Tensor] = None, *, frozen=None, return_fit: bool = False):
    """DESeq2's ``rlog`` (blind design): ridge NB GLM with one coefficient per sample, prior variance by weighted quantile
    matching, dispersions from the fitted trend (``models.deseq``).  ``frozen`` = a training set's ``RlogFit``: new samples are
    fitted against its intercepts, trend and prior variance."""
    from . import deseq as DQ
    if frozen is not None:
        r = DQ.rlog(counts, frozen=frozen)
    else:
        r = DQ.rlog(counts, DQ.estimate_dispersions(counts, sf, None, "parametric"))
    return (r.values, r) if return_fit else r.values


# ------------------------------------------------------------------------------------------------ classifiers
def _class_indicator(y: torch.Tensor, k: int) -> torch.Tensor:
    return torch.nn.functional.one_hot(y.long(), k).double()       # [n, k]


@dataclass
class PoissonLDA:
    """Witten's Poisson linear discriminant analysis; ``power=True`` gives PLDA2 (power-transformed counts)."""
    rho: float = 0.0
    beta: float = 1.0
    power: bool = False
    alpha_: float = 1.0
    d_: torch.Tensor = field(default=None, repr=False)
    g_: torch.Tensor = field(default=None, repr=False)
    prior_: torch.Tensor = field(default=None, repr=False)

    def _tx(self, c):
        return c.double() ** self.alpha_ if self.power else c.double()

    def fit(self, counts: torch.Tensor, y: torch.Tensor, n_classes: int = 2) -> "PoissonLDA":
        c = counts.double()
        if self.power:
            # choose the exponent that makes the counts look Poisson: chi-square goodness-of-fit closest to its df
            best, besta = None, 1.0
            for a in np.linspace(0.2, 1.0, 17):
                x = c ** a
                e = x.sum(1, keepdim=True) * x.sum(0, keepdim=True) / x.sum()
                stat = float((((x - e) ** 2) / e.clamp_min(1e-12)).sum())
                dfree = (x.shape[0] - 1) * (x.shape[1] - 1)
                if best is None or abs(stat - dfree) < best:
                    best, besta = abs(stat - dfree), float(a)
            self.alpha_ = besta
        x = self._tx(c)
        s = x.sum(0) / x.sum()                                      # total-count size factors
        g = x.sum(1)                                                # gene totals
        Y = _class_indicator(y, n_classes).to(x.device)
        a = x @ Y + self.beta                                       # [G, k] class sums
        b = g[:, None] * (s @ Y)[None, :] + self.beta
        d = a / b
        if self.rho > 0:
            dev = d - 1
            d = 1 + torch.sign(dev) * (dev.abs() - self.rho / torch.sqrt(b)).clamp_min(0)
        self.d_, self.g_ = d.clamp_min(1e-12), g
        self.prior_ = Y.mean(0)
        return self

    def scores(self, counts: torch.Tensor) -> torch.Tensor:
        x = self._tx(counts)
        s = x.sum(0) / self.g_.sum()
        if _native.on_device(x):
            G, n = x.shape
            k = self.d_.shape[1]
            out = torch.empty(1, n, k, dtype=torch.float64, device=x.device)
            P = _native.ptr
            rc = _native.cuda().cml_plda_scores(P(x.contiguous()), G, n, k, 1, P(torch.log(self.d_).contiguous()),
                                                P((self.g_[:, None] * self.d_).contiguous()), P(s.contiguous()),
                                                P(torch.log(self.prior_).reshape(1, k).contiguous()), P(out), _native.stream_ptr())
            _native.check(rc, "cml_plda_scores")
            return out[0]
        return x.t() @ torch.log(self.d_) - s[:, None] * (self.g_[:, None] * self.d_).sum(0)[None, :] + torch.log(self.prior_)[None, :]

    def predict(self, counts: torch.Tensor) -> torch.Tensor:
        return self.scores(counts).argmax(1)

    def selected_genes(self) -> torch.Tensor:
        return torch.nonzero((self.d_ != 1).any(1)).flatten()


@dataclass
class NBLDA:
    """Negative-binomial LDA with gene-wise moment dispersions (shrunk towards zero by ``shrink``)."""
    beta: float = 1.0
    shrink: float = 0.0
    d_: torch.Tensor = field(default=None, repr=False)
    g_: torch.Tensor = field(default=None, repr=False)
    phi_: torch.Tensor = field(default=None, repr=False)
    prior_: torch.Tensor = field(default=None, repr=False)

    def fit(self, counts: torch.Tensor, y: torch.Tensor, n_classes: int = 2) -> "NBLDA":
        x = counts.double()
        s = x.sum(0) / x.sum()
        g = x.sum(1)
        Y = _class_indicator(y, n_classes).to(x.device)
        self.d_ = ((x @ Y + self.beta) / (g[:, None] * (s @ Y)[None, :] + self.beta)).clamp_min(1e-12)
        mu = g[:, None] * s[None, :]
        phi = (((x - mu) ** 2 - mu).sum(1) / (mu * mu).sum(1).clamp_min(1e-12)).clamp_min(0)
        self.phi_ = phi * (1 - self.shrink)
        self.g_, self.prior_ = g, Y.mean(0)
        return self

    def scores(self, counts: torch.Tensor) -> torch.Tensor:
        x = counts.double()
        s = x.sum(0) / self.g_.sum()
        out = []
        for k in range(self.d_.shape[1]):
            m = s[None, :] * (self.g_ * self.d_[:, k])[:, None]              # [G, n] class-k means
            phi = self.phi_[:, None]
            nb = x * (torch.log(m) - torch.log1p(m * phi)) - torch.log1p(m * phi) / phi.clamp_min(1e-12)
            po = x * torch.log(m) - m
            out.append(torch.where(phi > 1e-10, nb, po).sum(0) + torch.log(self.prior_[k]))
        return torch.stack(out, 1)

    def predict(self, counts: torch.Tensor) -> torch.Tensor:
        return self.scores(counts).argmax(1)


@dataclass
class NearestShrunkenCentroids:
    """PAM / NSC on continuous data (genes x samples); ``weights`` (genes x samples precision weights from voom) turn it
    into voomNSC, and ``threshold = 0`` with weights into voomDLDA."""
    threshold: float = 0.0
    centroid_: torch.Tensor = field(default=None, repr=False)
    overall_: torch.Tensor = field(default=None, repr=False)
    sd_: torch.Tensor = field(default=None, repr=False)
    dprime_: torch.Tensor = field(default=None, repr=False)
    prior_: torch.Tensor = field(default=None, repr=False)

    def fit(self, x: torch.Tensor, y: torch.Tensor, n_classes: int = 2, weights: Optional[torch.Tensor] = None):
        x = x.double()
        n = x.shape[1]
        w = torch.ones_like(x) if weights is None else weights.double()
        Y = _class_indicator(y, n_classes).to(x.device)
        nk = Y.sum(0)
        wk = w @ Y
        cent = (w * x) @ Y / wk.clamp_min(1e-12)
        overall = (w * x).sum(1) / w.sum(1)
        resid = x - cent @ Y.t()
        s = torch.sqrt((w * resid * resid).sum(1) / w.mean(1) / (n - n_classes))
        s0 = s.median()
        mk = torch.sqrt(1.0 / nk - 1.0 / n)
        d = (cent - overall[:, None]) / (mk[None, :] * (s + s0)[:, None])
        dp = torch.sign(d) * (d.abs() - self.threshold).clamp_min(0)
        self.centroid_ = overall[:, None] + mk[None, :] * (s + s0)[:, None] * dp
        self.overall_, self.sd_, self.dprime_, self.prior_ = overall, s + s0, dp, nk / n
        return self

    def scores(self, x: torch.Tensor, weights: Optional[torch.Tensor] = None) -> torch.Tensor:
        x = x.double()
        w = torch.ones_like(x) if weights is None else weights.double() / weights.double().mean(1, keepdim=True)
        out = []
        for k in range(self.centroid_.shape[1]):
            z = (x - self.centroid_[:, k:k + 1]) / self.sd_[:, None]
            out.append(-(w * z * z).sum(0) + 2 * torch.log(self.prior_[k]))
        return torch.stack(out, 1)

    def predict(self, x: torch.Tensor, weights: Optional[torch.Tensor] = None) -> torch.Tensor:
        return self.scores(x, weights).argmax(1)

    def selected_genes(self) -> torch.Tensor:
        return torch.nonzero((self.dprime_ != 0).any(1)).flatten()


# ------------------------------------------------------------------------------------------------ the MLSeq front-end
def _voom_transform(counts: torch.Tensor, y: torch.Tensor, n_classes: int):
    """log2-cpm and voom precision weights for a class-means design (oracle lowess on CPU, kernels on GPU)."""
    from ..oracle import limma as O
    c = counts.double().cpu().numpy()
    lib = c.sum(0)
    design = np.eye(n_classes)[y.cpu().numpy().astype(int)]
    e, w, _ = O.voom(c, design, lib)
    return torch.as_tensor(e, device=counts.device), torch.as_tensor(w, device=counts.device)


METHODS = ("PLDA", "PLDA2", "NBLDA", "pam", "voomNSC", "voomDLDA")


@dataclass
class QDA:
    """Quadratic discriminant analysis with the class covariances shrunk towards their diagonal (``shrink`` = 0 is
    plain QDA; needed as soon as genes outnumber the samples of a class).  The reference only stubs this
    (BuieRProj/KM_WHATEVER.Rmd:375-386 prepares the frame and stops)."""
    shrink: float = 0.1
    means_: torch.Tensor = field(default=None, repr=False)
    chol_: list = field(default_factory=list, repr=False)
    logdet_: torch.Tensor = field(default=None, repr=False)
    logprior_: torch.Tensor = field(default=None, repr=False)

    def fit(self, x: torch.Tensor, y: torch.Tensor, n_classes: int = 2) -> "QDA":
        x = x.double()
        self.means_, self.chol_, ld, lp = [], [], [], []
        for c in range(n_classes):
            xc = x[y.to(x.device) == c]
            mu = xc.mean(0)
            d = xc - mu
            cov = d.t() @ d / max(xc.shape[0] - 1, 1)
            diag = torch.diag(torch.diagonal(cov).clamp_min(1e-8))
            cov = (1 - self.shrink) * cov + self.shrink * diag + 1e-8 * torch.eye(x.shape[1], dtype=x.dtype, device=x.device)
            L = torch.linalg.cholesky(cov)
            self.means_.append(mu)
            self.chol_.append(L)
            ld.append(2 * torch.log(torch.diagonal(L)).sum())
            lp.append(np.log(max(xc.shape[0], 1) / x.shape[0]))
        self.logdet_ = torch.stack(ld)
        self.logprior_ = torch.tensor(lp, dtype=x.dtype, device=x.device)
        return self

    def decision_function(self, x: torch.Tensor) -> torch.Tensor:
        x = x.double()
        sc = []
        for mu, L, ld, lp in zip(self.means_, self.chol_, self.logdet_, self.logprior_):
            z = torch.linalg.solve_triangular(L, (x - mu).t(), upper=False)
            sc.append(-0.5 * (z * z).sum(0) - 0.5 * ld + lp)
        return torch.stack(sc, 1)

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return self.decision_function(x).argmax(1)

    def selected_genes(self) -> torch.Tensor:
        return torch.arange(self.means_[0].numel())


def _svm_radial(cost, seed):
    from .svm import SVC
    return SVC(kernel="radial", cost=float(cost))


def _logitboost(n_iter, seed):
    # caTools::LogitBoost = Newton boosting of decision stumps on the logistic loss
    from .trees import GradientBoosting
    return GradientBoosting(n_estimators=int(n_iter), max_depth=1, learning_rate=0.5, reg_lambda=0.0, min_child_weight=1e-3,
                            mode="xgboost", seed=seed)


def _blackboost(depth, seed):
    # mboost::blackboost = first-order boosting of shallow regression trees, nu = 0.1
    from .trees import GradientBoosting
    return GradientBoosting(n_estimators=100, max_depth=int(depth), learning_rate=0.1, mode="sklearn", seed=seed)


def _deepboost(depth, seed):
    # deepboost: boosted trees with a complexity penalty that grows with depth -> split penalty gamma, L2 leaf penalty
    from .trees import GradientBoosting
    return GradientBoosting(n_estimators=50, max_depth=int(depth), learning_rate=0.3, reg_lambda=1.0, gamma=0.05 * int(depth),
                            mode="xgboost", seed=seed)


_TRANSFORMED = {"svmRadial": _svm_radial, "LogitBoost": _logitboost, "blackboost": _blackboost, "deepboost": _deepboost,
                "qda": lambda shrink, seed: QDA(shrink=float(shrink))}


class _TransformedModel:
    """A learner trained on VST / rlog values; ``predict`` takes raw counts (genes x samples) like the count models."""

    def __init__(self, model, pre, train_fit):
        self.model, self.pre, self.train_fit = model, pre, train_fit

    def _transform(self, counts):
        # new samples against the TRAINING set's geometric means, dispersion function, rlog intercepts and prior variance
        # (MLSeq: estimateSizeFactors(geoMeans = ), dispersionFunction<-, rlog(intercept = , betaPriorVar = ))
        t = deseq_vst(counts, fit=self.train_fit) if self.pre == "deseq-vst" else deseq_rlog(counts, frozen=self.train_fit)
        return t.float().t().contiguous()

    def predict(self, counts: torch.Tensor) -> torch.Tensor:
        return self.model.predict(self._transform(counts))

    def selected_genes(self) -> torch.Tensor:
        m = self.model
        if hasattr(m, "importance_") and m.importance_ is not None:
            return torch.nonzero(m.importance_["gain"] > 0).flatten().cpu()
        if hasattr(m, "selected_genes"):
            return m.selected_genes()
        return torch.arange(m.x_.shape[1])


def classify(counts_train: torch.Tensor, y_train: torch.Tensor, method: str = "PLDA", tune: Optional[Sequence[float]] = None,
             folds: int = 5, repeats: int = 10, seed: int = 2128, pre: str = "deseq-vst", n_classes: int = 2):
    """MLSeq-style ``classify``: tune the method's shrinkage parameter by repeated k-fold CV (5 x 10 in the reference,
    VikasP/AML.R:241-245), refit on all training samples.  Returns (model, cv table {param: accuracy})."""
    from ..ops.replicate import replicate_weights
    n = counts_train.shape[1]
    y_train = y_train.to(counts_train.device)
    if tune is None:
        tune = {"PLDA": np.linspace(0, 15, 10), "PLDA2": np.linspace(0, 15, 10), "NBLDA": [0.0],
                "pam": np.linspace(0, 4, 10), "voomNSC": np.linspace(0, 4, 10), "voomDLDA": [0.0],
                "svmRadial": [2.0 ** k for k in range(-2, 8)],          # caret: C = 2^((1:tuneLength) - 3)
                "LogitBoost": [11, 21, 31],                             # caret: nIter
                "blackboost": [1, 2, 3],                                # caret: maxdepth (mstop fixed below)
                "deepboost": [1, 2, 3],                                 # caret: tree_depth
                "qda": [0.0, 0.1, 0.5]}[method]                         # shrinkage towards the diagonal
    masks = replicate_weights(folds * repeats, n, mode="kfold", k=folds, seed=seed, device="cpu")[:, :n] > 0     # True = train

    def make(param):
        if method == "PLDA":
            return PoissonLDA(rho=float(param))
        if method == "PLDA2":
            return PoissonLDA(rho=float(param), power=True)
        if method == "NBLDA":
            return NBLDA()
        if method in _TRANSFORMED:
            return _TRANSFORMED[method](param, seed)
        return NearestShrunkenCentroids(threshold=float(param))

    transformed = train_fit = None
    if method in _TRANSFORMED or method == "pam":                 # learners on the VST / rlog matrix
        transformed, train_fit = deseq_vst(counts_train, return_fit=True) if pre == "deseq-vst" else deseq_rlog(counts_train, return_fit=True)
        transformed = transformed.float() if method in _TRANSFORMED else transformed

    def fit_predict(param, tr, te):
        ctr, cte, ytr = counts_train[:, tr], counts_train[:, te], y_train[tr]
        mdl = make(param)
        if method in _TRANSFORMED:
            return mdl.fit(transformed[:, tr].t().contiguous(), ytr).predict(transformed[:, te].t().contiguous())
        if method in ("PLDA", "PLDA2", "NBLDA"):
            return mdl.fit(ctr, ytr, n_classes).predict(cte)
        if method == "pam":
            return mdl.fit(transformed[:, tr], ytr, n_classes).predict(transformed[:, te])
        e, w = _voom_transform(ctr, ytr, n_classes)
        mdl.fit(e, ytr, n_classes, weights=w)
        lib = cte.double().sum(0)
        ete = torch.log2((cte.double() + 0.5) / (lib + 1)[None, :] * 1e6)
        return mdl.predict(ete)

    table: Dict[float, float] = {}
    for param in tune:
        acc = []
        for r in range(masks.shape[0]):
            tr = torch.nonzero(masks[r]).flatten().to(counts_train.device)
            te = torch.nonzero(~masks[r]).flatten().to(counts_train.device)
            if te.numel() == 0 or len(torch.unique(y_train[tr])) < n_classes:
                continue
            acc.append(float((fit_predict(param, tr, te) == y_train[te].long()).float().mean()))
        table[float(param)] = float(np.mean(acc)) if acc else float("nan")
    best = max(table, key=lambda k: (table[k], k))
    mdl = make(best)
    if method in _TRANSFORMED:
        mdl = _TransformedModel(mdl.fit(transformed.t().contiguous(), y_train), pre, train_fit)
    elif method in ("PLDA", "PLDA2", "NBLDA"):
        mdl.fit(counts_train, y_train, n_classes)
    elif method == "pam":
        mdl.fit(transformed, y_train, n_classes)
        mdl.train_fit_, mdl.pre_ = train_fit, pre                 # new samples go through the training set's transform
    else:
        e, w = _voom_transform(counts_train, y_train, n_classes)
        mdl.fit(e, y_train, n_classes, weights=w)
    mdl.best_param_ = best
    return mdl, table


def selected_genes(model, gene_ids: Sequence[str]) -> List[str]:
    """MLSeq ``selectedGenes``: the genes a sparse classifier kept (VikasP/AML.R:279-284)."""
    return [gene_ids[i] for i in model.selected_genes().tolist()]
