"""Logistic lasso / elastic-net with cross-validated lambda - the reference's headline model.

``run_lasso`` restates ``runLasso`` (composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:68-123) and
``glm.binom`` (JSmith_code/Differential_Expression_and_Lasso.Rmd:302-391): train path, LOOCV for
lambda with the class-error measure, held-out confusion matrix, full-data refit, non-zero coefficients.
``exclusion_reps`` restates the 15-replicate loop that removes previously selected genes
(cml_targetaml_seanalysis.Rmd:728-773).  ``stability_selection`` is the north-star extension
(bootstrap x alpha grid).  All replicate fits of one call (train fit + every CV fold + refit) are one
batched kernel launch.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import _native
from ..data import Cohort
from ..ops import lasso as K
from ..ops import metrics as M


@dataclass
class CVResult:
    lam: np.ndarray
    cvm: np.ndarray
    cvsd: np.ndarray
    nzero: np.ndarray
    lambda_min: float
    lambda_1se: float
    index_min: int
    index_1se: int
    measure: str
    foldid: np.ndarray


@dataclass
class LassoResult:
    """Mirror of the reference's ``resultslist`` (cml_targetaml_seanalysis.Rmd:117-121)."""

    training_set: List[str]
    testing_set: List[str]
    gene_ids: List[str]
    lam: np.ndarray
    train_beta: np.ndarray          # [L, p]
    train_dev_ratio: np.ndarray
    train_npasses: int
    cv: CVResult
    confusion: np.ndarray           # table(pred, truth)
    test_error: float
    test_auc: float
    test_logloss: float
    final_beta: np.ndarray          # [L, p]
    final_npasses: int
    nonzero_coef: Dict[str, float]  # gene id -> coefficient at lambda.min of the full-data refit
    seed: int
    jerr: int = 0
    report: Optional[M.BinaryReport] = None


def make_foldid(n: int, nfolds: int, seed: int = 2019) -> np.ndarray:
    """``sample(rep(seq(nfolds), length = n))`` - balanced random fold labels (cv.glmnet's default)."""
    rng = np.random.default_rng(seed)
    base = np.arange(n) % nfolds
    return base[rng.permutation(n)]


def _cv_weights(n: int, foldid: np.ndarray, base_w: np.ndarray) -> np.ndarray:
    nf = int(foldid.max()) + 1
    w = np.repeat(base_w[None, :], nf, axis=0)
    for f in range(nf):
        w[f, foldid == f] = 0.0
    return w


def _getmin(lam, cvm, cvsd):
    cvmin = np.nanmin(cvm)
    idmin = cvm <= cvmin
    lambda_min = lam[idmin].max()
    imin = int(np.flatnonzero(lam == lambda_min)[0])
    semin = cvm[imin] + cvsd[imin]
    id1 = cvm <= semin
    lambda_1se = lam[id1].max()
    i1 = int(np.flatnonzero(lam == lambda_1se)[0])
    return lambda_min, lambda_1se, imin, i1


def cv_from_eta(eta_folds: torch.Tensor, y: torch.Tensor, foldid: np.ndarray, lam: np.ndarray, nzero: np.ndarray,
                weights: Optional[np.ndarray] = None, measure: str = "class") -> CVResult:
    """Turn the held-out linear predictors of every fold fit into cv.glmnet's ``cvm/cvsd/lambda.min``.

    eta_folds : [nfolds, L, n] (fold f's fit evaluated on all samples).  Class error follows cv.lognet:
    a prediction of exactly 0.5 counts as class 0; with < 3 samples per fold (LOOCV) the error is
    averaged per observation (``grouped = FALSE``)."""
    nf, L, n = eta_folds.shape
    fid = torch.as_tensor(foldid, device=eta_folds.device)
    # pick, for each sample, the predictor from the fold that left it out -> [L, n]
    pred = eta_folds[fid, :, torch.arange(n, device=eta_folds.device)].t()
    yy = y.to(pred.dtype)[None, :]
    if measure == "class":
        raw = torch.where(pred > 0, 1.0 - yy, yy)          # link > 0 <=> prob > .5
    elif measure == "deviance":
        p = torch.sigmoid(pred).clamp(1e-5, 1 - 1e-5)
        raw = -2 * (yy * p.log() + (1 - yy) * (1 - p).log())
    else:
        raise ValueError(measure)
    raw = raw.double().cpu().numpy()                        # [L, n]
    w = np.ones(n) if weights is None else np.asarray(weights, float)
    counts = np.bincount(foldid, minlength=nf)
    grouped = counts.min() >= 3
    if grouped:
        # per-fold means, then weighted mean / sd over folds
        fm = np.stack([np.average(raw[:, foldid == f], axis=1, weights=w[foldid == f]) for f in range(nf)], 1)
        fw = np.array([w[foldid == f].sum() for f in range(nf)])
        cvm = np.average(fm, axis=1, weights=fw)
        cvsd = np.sqrt(np.average((fm - cvm[:, None]) ** 2, axis=1, weights=fw) / (nf - 1))
    else:
        cvm = np.average(raw, axis=1, weights=w)
        cvsd = np.sqrt(np.average((raw - cvm[:, None]) ** 2, axis=1, weights=w) / (n - 1))
    lmin, l1se, imin, i1 = _getmin(lam, cvm, cvsd)
    return CVResult(lam, cvm, cvsd, nzero, float(lmin), float(l1se), imin, i1, measure, foldid)


class _side_stream:
    """Run the enclosed launches on their own CUDA stream so that independent path batches overlap on the GPU
    (a single fit occupies one warp, so the train fit, the CV batch and the refit would otherwise queue up behind
    each other for ~0.13 s each).  No-op for CPU tensors."""

    def __init__(self, enabled: bool):
        self.enabled = enabled and torch.cuda.is_available()
        self.stream = torch.cuda.Stream() if self.enabled else None
        self.ctx = None

    def __enter__(self):
        if self.enabled:
            self.stream.wait_stream(torch.cuda.current_stream())
            self.ctx = torch.cuda.stream(self.stream)
            self.ctx.__enter__()
        return self

    def __exit__(self, *a):
        if self.enabled:
            self.ctx.__exit__(*a)

    def join(self):
        if self.enabled:
            torch.cuda.current_stream().wait_stream(self.stream)


def fit_cv(x: torch.Tensor, y: torch.Tensor, lam: Optional[np.ndarray] = None, *, nfolds: Optional[int] = None,
           foldid: Optional[np.ndarray] = None, alpha: float = 1.0, intercept: bool = False,
           measure: str = "class", seed: int = 2019, feature_mask: Optional[torch.Tensor] = None,
           cv_modified_newton: bool = True, extra_weights: Optional[np.ndarray] = None, nx: Optional[int] = None):
    """Path fit on all of ``x`` plus a CV fit per fold, as ONE replicate batch.

    Returns (PathBatch, CVResult).  Batch layout: replicate 0 = fit on all samples (Newton, like
    ``glmnet()``); replicate 1 = the same with the CV solver variant (it is ``cv.fit$glmnet.fit`` and
    supplies ``nzero``); replicates 2.. = fold fits.  ``nfolds=None`` means leave-one-out, as in the
    reference (``nfolds = length(train.names)``, cml_targetaml_seanalysis.Rmd:100)."""
    n = x.shape[0]
    lam = K.default_lambda_grid() if lam is None else np.asarray(lam, float)
    if foldid is None:
        nfolds = n if nfolds is None else nfolds
        foldid = np.arange(n) if nfolds == n else make_foldid(n, nfolds, seed)
    foldid = np.asarray(foldid)
    base_w = np.ones(n)
    wf = _cv_weights(n, foldid, base_w)
    dev = x.device
    if cv_modified_newton:
        # two launches: solver variant is a launch-level switch
        w0 = torch.as_tensor(base_w[None, :], device=dev)
        packed = K.pack_design(x) if _native.on_device(x) else None
        wcv = torch.as_tensor(np.concatenate([base_w[None, :], wf], 0), device=dev)
        s_full, s_cv = _side_stream(x.is_cuda), _side_stream(x.is_cuda)
        with s_full:
            full = K.lognet_path(x, y, w0, lam, alpha=alpha, intercept=intercept, modified_newton=False,
                                 feature_mask=feature_mask, packed=packed, nx=nx)
        with s_cv:
            folds = K.lognet_path(x, y, wcv, lam, alpha=alpha, intercept=intercept, modified_newton=True,
                                  feature_mask=feature_mask, packed=packed, nx=nx)
        s_full.join()
        s_cv.join()
        nzero = folds.df()[0].cpu().numpy()
        eta_f = folds.eta[1:]
    else:
        wall = torch.as_tensor(np.concatenate([base_w[None, :], wf], 0), device=dev)
        folds = K.lognet_path(x, y, wall, lam, alpha=alpha, intercept=intercept, modified_newton=False,
                              feature_mask=feature_mask, nx=nx)
        full = folds
        nzero = folds.df()[0].cpu().numpy()
        eta_f = folds.eta[1:]
    cv = cv_from_eta(eta_f, y, foldid, lam, nzero, measure=measure)
    return full, folds, cv


def run_lasso(cohort: Cohort, seed: int = 2019, lam: Optional[np.ndarray] = None, alpha: float = 1.0,
              nfolds: Optional[int] = None, label_col: str = "deg.risk", split_col: str = "exptset.seahack",
              device: Optional[torch.device] = None) -> LassoResult:
    """The reference's ``runLasso(seset, seed)`` on a :class:`Cohort`."""
    lam = K.default_lambda_grid() if lam is None else np.asarray(lam, float)
    flag = cohort.col_data[split_col].astype(str).to_numpy()
    tr_idx, te_idx = np.flatnonzero(flag == "train"), np.flatnonzero(flag == "test")
    x_all = cohort.design(torch.float64)
    y_all = cohort.labels(label_col).to(torch.float64)
    if device is not None:
        x_all, y_all = x_all.to(device), y_all.to(device)
    if _native.on_device(x_all):
        x_all, y_all = x_all.float(), y_all.float()
    xtr, ytr = x_all[tr_idx], y_all[tr_idx]
    xte, yte = x_all[te_idx], y_all[te_idx]

    # the full-data refit does not depend on the CV result: start it first, on its own stream
    ones = torch.ones(1, x_all.shape[0], dtype=torch.float64, device=x_all.device)
    s_final = _side_stream(x_all.is_cuda)
    with s_final:
        final = K.lognet_path(x_all, y_all, ones, lam, alpha=alpha, intercept=False, modified_newton=False)
    full, folds, cv = fit_cv(xtr, ytr, lam, nfolds=nfolds, alpha=alpha, seed=seed)
    s_final.join()
    beta_tr = full.dense(0)
    eta_te = xte.to(beta_tr.dtype) @ beta_tr[cv.index_min] + full.a0[0, cv.index_min]
    pred = (eta_te > 0).long()
    rep = M.binary_report(pred, yte)
    prob = torch.sigmoid(eta_te)

    beta_f = final.dense(0)
    coef = beta_f[cv.index_min].double().cpu().numpy()
    nz = np.flatnonzero(coef != 0)
    genes = cohort.gene_ids
    return LassoResult(
        [cohort.sample_ids[i] for i in tr_idx], [cohort.sample_ids[i] for i in te_idx], list(genes), lam,
        beta_tr.double().cpu().numpy(), full.dev_ratio[0].double().cpu().numpy(), int(full.npasses[0]), cv,
        rep.confusion.cpu().numpy(), rep.error, M.auc(eta_te, yte), M.log_loss(yte, prob),
        beta_f.double().cpu().numpy(), int(final.npasses[0]), {genes[i]: float(coef[i]) for i in nz}, seed,
        int(full.jerr[0]) or int(final.jerr[0]), rep)


def exclusion_reps(cohort: Cohort, nreps: int = 15, seed: int = 2019, **kw) -> List[LassoResult]:
    """Iterative exclusion: refit after removing every gene selected so far
    (cml_targetaml_seanalysis.Rmd:728-773; the reference ran 15 replicates, 558 genes excluded in total)."""
    results: List[LassoResult] = []
    current = cohort
    for _ in range(nreps):
        res = run_lasso(current, seed=seed, **kw)
        results.append(res)
        if not res.nonzero_coef:
            break
        current = current.exclude_genes(list(res.nonzero_coef))
    return results


def stability_selection(x: torch.Tensor, y: torch.Tensor, lam: np.ndarray, *, n_boot: int = 1000,
                        alphas: Sequence[float] = (1.0,), subsample: Optional[float] = None, seed: int = 0,
                        batch: int = 4096, nx: Optional[int] = None) -> Dict[str, torch.Tensor]:
    """Selection frequency of every gene over bootstrap (or subsample) replicates x alpha grid x lambda grid.

    The north-star configuration ("1000 bootstrap replicates x elastic-net alpha grid", BASELINE.json
    config 5).  Bootstrap multiplicities enter as integer observation weights, so every replicate is one
    more row of the weight matrix in the same launch.  Returns ``freq[alpha, L, p]`` and ``max_freq[p]``."""
    from ..ops.replicate import replicate_weights
    n, p = x.shape
    dev = x.device
    if _native.on_device(x):
        packed = K.pack_design(x)
        n_pad = packed[1]
    else:
        packed, n_pad = None, n
    L = len(lam)
    freq = torch.zeros(len(alphas), L, p, dtype=torch.float32, device=dev)
    mode = "bootstrap" if subsample is None else "subsample"
    k = 0 if subsample is None else max(2, int(round(subsample * n)))
    for ai, alpha in enumerate(alphas):
        done = 0
        while done < n_boot:
            nb = min(batch, n_boot - done)
            w = replicate_weights(nb, n, mode=mode, k=k, seed=seed + 7919 * ai, device=dev, n_pad=n_pad, rep0=done)
            K.lognet_path(x, y, w[:, :n], lam, alpha=alpha, want_eta=False, packed=packed, nx=nx, freq_out=freq[ai],
                          padded_weights=w if _native.on_device(x) else None, want_coef=False)
            done += nb
        freq[ai] /= n_boot
    return {"freq": freq, "max_freq": freq.amax(dim=(0, 1)), "alphas": torch.tensor(list(alphas)), "lam": torch.as_tensor(lam)}


def l1_logistic(x: torch.Tensor, y: torch.Tensor, C: float = 1.0, intercept: bool = True, n_path: int = 30, thresh: float = 1e-9,
                positive_only: bool = False) -> Dict[str, torch.Tensor]:
    """L1-penalised logistic regression in scikit-learn's parametrisation,
    ``LogisticRegression(penalty='l1', solver='saga', C=C)`` as run on all 21,404 genes in the walkthrough
    (scripts/model_walkthrough.ipynb raw lines 1471-1476 and 3886-3891, converged with ``max_iter=10000``).

    sklearn minimises ``|w|_1 + C * sum_i logloss_i``; dividing by ``C n`` gives glmnet's objective with
    ``lambda = 1 / (C n)``, an unpenalised intercept and no standardisation - so the deterministic path solver reaches the
    optimum SAGA converges to, warm-started down a short geometric path.  ``positive_only`` reproduces the notebook's
    selection rule ``coef_ > 0`` (lines 1488-1489, 3914), which silently drops negative coefficients."""
    n = x.shape[0]
    lam_target = 1.0 / (C * n)
    xd = x.double() if not _native.on_device(x) else x.float()
    yd = y.to(xd.dtype)
    r0 = yd - yd.mean() if intercept else yd - 0.5
    lam_max = float((xd.t() @ r0).abs().max() / n)
    lam = np.geomspace(max(lam_max, lam_target * 1.0001), lam_target, n_path)
    pb = K.lognet_path(x, y, torch.ones(1, n, dtype=xd.dtype, device=x.device), lam, intercept=intercept, thresh=thresh,
                       nx=min(x.shape[1], 4096), want_eta=False)
    coef = pb.dense(0)[-1]
    sel = coef > 0 if positive_only else coef != 0
    return {"coef": coef, "intercept": pb.a0[0, -1], "selected": torch.nonzero(sel).flatten(), "lambda": lam_target,
            "jerr": pb.jerr[0], "npasses": pb.npasses[0]}
