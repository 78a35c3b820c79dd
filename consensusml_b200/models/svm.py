"""Support vector classification (linear / RBF) with the reference's ``runSVM`` protocol.

``run_svm`` restates ``runSVM(seed, kerneltype, trainset, trainclasses, testset, testclasses, weightfilt)``
(composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:126-210): e1071 C-classification with ``scale = TRUE``,
cost 1, gamma = 1/p; weight vector ``t(coefs) %*% SV``; optional refit on the top ``weightfilt`` fraction of
|weight| columns; decision values, ROC, test precision / recall.  ``weight_quantile_tails`` restates the
quantile gene lists of BuieRProj/KM_WHATEVER.Rmd:315-362.

CUDA path: z-scoring is fused into the bf16 conversion, the kernel matrix is one tcgen05 GEMM with the RBF
exponent in its epilogue (ops/gemm.py), the dual is solved by ``cml_svm_smo`` (csrc/svm.cu, libsvm's SMO with
second-order working-set selection), one CTA per (C, fold) problem.  CPU path: NumPy restatement of the
same solver (``smo_numpy``).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import _native
from ..ops import gemm as G
from ..ops import metrics as M

_I, _LL, _F, _P = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p
_native.register_signature("cml_svm_smo", [_I, _I, _P, _LL, _LL, _P, _LL, _P, _LL, _F, _I, _P, _P, _P, _P, _P])
_native.register_signature("cml_svm_smo_warm", [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_void_p,
                                                ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_float, ctypes.c_int, ctypes.c_void_p,
                                                ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                ctypes.c_void_p])
_native.register_signature("cml_svm_decision", [_P, _LL, _LL, _I, _I, _P, _P, _I, _P, _P])


# ------------------------------------------------------------------------------------------------ solver
def smo_numpy(K: np.ndarray, y: np.ndarray, cbound: np.ndarray, eps: float = 1e-3, max_iter: int = 1000000,
              alpha0: Optional[np.ndarray] = None, g0: Optional[np.ndarray] = None, return_gradient: bool = False):
    """libsvm's Solver (WSS2) in float64.  Returns (alpha, rho, iterations) [+ gradient].  ``alpha0`` / ``g0`` warm-start it
    (sub-problem of a working set: ``g0`` is the global gradient Q alpha - e restricted to these variables)."""
    n = K.shape[0]
    G_ = -np.ones(n) if g0 is None else np.array(g0, dtype=np.float64)
    a = np.zeros(n) if alpha0 is None else np.array(alpha0, dtype=np.float64)
    kd = np.diag(K).copy()
    it = 0
    for it in range(max_iter):
        inp = cbound > 0
        up = inp & (((y > 0) & (a < cbound)) | ((y < 0) & (a > 0)))
        low = inp & (((y > 0) & (a > 0)) | ((y < 0) & (a < cbound)))
        if not up.any():
            break
        v = np.where(up, -y * G_, -np.inf)
        i = int(np.argmax(v))
        gmax = v[i]
        ytg = y * G_
        g2 = np.max(np.where(low, ytg, -np.inf)) if low.any() else -np.inf
        bq = gmax + ytg
        aq = kd[i] + kd - 2 * K[i]
        aq = np.where(aq <= 0, 1e-12, aq)
        cand = low & (bq > 0)
        if gmax + g2 < eps or not cand.any():
            break
        j = int(np.argmax(np.where(cand, bq * bq / aq, -np.inf)))
        yi, yj, Ci, Cj = y[i], y[j], cbound[i], cbound[j]
        ai, aj = a[i], a[j]
        quad = kd[i] + kd[j] - 2 * K[i, j]
        if quad <= 0:
            quad = 1e-12
        if yi != yj:
            d = (-G_[i] - G_[j]) / quad
            diff = ai - aj
            ai += d
            aj += d
            if diff > 0:
                if aj < 0:
                    aj, ai = 0.0, diff
            elif ai < 0:
                ai, aj = 0.0, -diff
            if diff > Ci - Cj:
                if ai > Ci:
                    ai, aj = Ci, Ci - diff
            elif aj > Cj:
                aj, ai = Cj, Cj + diff
        else:
            d = (G_[i] - G_[j]) / quad
            s = ai + aj
            ai -= d
            aj += d
            if s > Ci:
                if ai > Ci:
                    ai, aj = Ci, s - Ci
            elif aj < 0:
                aj, ai = 0.0, s
            if s > Cj:
                if aj > Cj:
                    aj, ai = Cj, s - Cj
            elif ai < 0:
                ai, aj = 0.0, s
        G_ += y * (K[i] * ((ai - a[i]) * yi) + K[j] * ((aj - a[j]) * yj))
        a[i], a[j] = ai, aj
    inp = cbound > 0
    yg = y * G_
    at_ub = inp & (a >= cbound)
    at_lb = inp & (a <= 0)
    free = inp & ~at_ub & ~at_lb
    if free.any():
        rho = yg[free].mean()
    else:
        ub_set = (at_ub & (y < 0)) | (at_lb & (y > 0))
        lb_set = (at_ub & (y > 0)) | (at_lb & (y < 0))
        ub = yg[ub_set].min() if ub_set.any() else np.inf
        lb = yg[lb_set].max() if lb_set.any() else -np.inf
        rho = 0.5 * (ub + lb)
    if return_gradient:
        return a, float(rho), it, G_
    return a, float(rho), it


def solve_dual(K: torch.Tensor, y_pm: torch.Tensor, cbound: torch.Tensor, eps: float = 1e-3, max_iter: int = 200000):
    """Solve B dual problems.  K: [n, n] (shared) or [B, n, n]; y_pm: [n] or [B, n] in {-1, +1};
    cbound: [B, n] per-sample cost (0 removes the sample).  Returns (alpha [B, n], rho [B], iters [B])."""
    B, n = cbound.shape
    if not _native.on_device(K):
        Kn = K.double().numpy()
        yn = y_pm.double().numpy()
        out = [smo_numpy(Kn if Kn.ndim == 2 else Kn[b], yn if yn.ndim == 1 else yn[b], cbound[b].double().numpy(), eps, max_iter)
               for b in range(B)]
        return (torch.tensor(np.stack([o[0] for o in out])), torch.tensor([o[1] for o in out]),
                torch.tensor([o[2] for o in out], dtype=torch.int32))
    dev = K.device
    K = K.float()
    assert K.stride(-1) == 1
    ldk = K.stride(-2)
    k_stride = K.stride(0) if K.dim() == 3 else 0
    y_pm = y_pm.float().contiguous()
    y_stride = n if y_pm.dim() == 2 else 0
    cb = cbound.float().contiguous()
    alpha = torch.empty(B, n, dtype=torch.float32, device=dev)
    rho = torch.empty(B, dtype=torch.float32, device=dev)
    iters = torch.empty(B, dtype=torch.int32, device=dev)
    obj = torch.empty(B, dtype=torch.float32, device=dev)
    rc = _native.cuda().cml_svm_smo(B, n, ctypes.c_void_p(K.data_ptr()), k_stride, ldk, _native.ptr(y_pm), y_stride,
                                    _native.ptr(cb), n, eps, max_iter, _native.ptr(alpha), _native.ptr(rho),
                                    _native.ptr(iters), _native.ptr(obj), _native.stream_ptr())
    _native.check(rc, "cml_svm_smo")
    return alpha, rho, iters


# ------------------------------------------------------------------------------------------------ model
@dataclass
class SVC:
    """C-support vector classifier with e1071's defaults (``scale=TRUE``, ``cost=1``, ``gamma=1/p``)."""
    kernel: str = "linear"            # 'linear' | 'radial'
    cost: float = 1.0
    gamma: Optional[float] = None
    scale: bool = True
    tol: float = 1e-3

    center_: torch.Tensor = field(default=None, repr=False)
    scale_: torch.Tensor = field(default=None, repr=False)
    x_: torch.Tensor = field(default=None, repr=False)            # scaled training data
    coef_: torch.Tensor = field(default=None, repr=False)         # alpha * y  [n]
    rho_: float = 0.0
    iters_: int = 0

    def _prep(self, x):
        if self.scale:
            return (x - self.center_) / self.scale_
        return x

    def _kernel(self, a, b=None):
        p = a.shape[1]
        if self.kernel == "linear":
            return G.linear_kernel(a, b)
        if self.kernel == "radial":
            return G.rbf_kernel(a, b, self.gamma if self.gamma is not None else 1.0 / p)
        raise ValueError(self.kernel)

    def fit(self, x: torch.Tensor, y: torch.Tensor, sample_mask: Optional[torch.Tensor] = None) -> "SVC":
        x = x.float()
        if self.scale:
            self.center_ = x.mean(0)
            sd = x.std(0, unbiased=True)
            self.scale_ = torch.where(sd > 0, sd, torch.ones_like(sd))
        self.x_ = self._prep(x).contiguous()
        K = self._kernel(self.x_)
        ypm = (2 * y.float() - 1).to(x.device)
        cb = torch.full((1, x.shape[0]), float(self.cost), device=x.device)
        if sample_mask is not None:
            cb = cb * sample_mask.float()[None, :]
        alpha, rho, iters = solve_dual(K.contiguous(), ypm, cb, self.tol)
        self.coef_ = (alpha[0].to(ypm.dtype) * ypm).float()
        self.rho_ = float(rho[0])
        self.iters_ = int(iters[0])
        return self

    @property
    def support_(self) -> torch.Tensor:
        return torch.nonzero(self.coef_ != 0).flatten()

    def decision_function(self, x: torch.Tensor) -> torch.Tensor:
        """Positive = class 1 (the reference's ``decision.values``, up to e1071's level-order sign)."""
        Kt = self._kernel(self._prep(x.float()).contiguous(), self.x_)
        return Kt.float() @ self.coef_ - self.rho_

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return (self.decision_function(x) > 0).long()

    def weights(self) -> torch.Tensor:
        """``t(svm_model$coefs) %*% svm_model$SV`` - the primal weight vector in scaled feature space
        (exact for the linear kernel; for 'radial' it is the same linear combination of support vectors the
        reference computes, cml_targetaml_seanalysis.Rmd:155)."""
        return self.coef_ @ self.x_


@dataclass
class SVMResult:
    """Mirror of ``runSVM``'s return list."""
    options: List[str]
    model: SVC
    weightsvect: torch.Tensor
    columns: torch.Tensor               # feature indices the final model was fitted on
    decision_train: torch.Tensor
    decision_test: torch.Tensor
    pred_test: torch.Tensor
    roc: tuple
    auc: float
    precision_test: float
    recall_test: float
    test_error: float


def run_svm(xtr: torch.Tensor, ytr: torch.Tensor, xte: torch.Tensor, yte: torch.Tensor, kerneltype: str = "linear",
            weightfilt: Optional[float] = None, seed: int = 50) -> SVMResult:
    opts = []
    cols = torch.arange(xtr.shape[1], device=xtr.device)
    model = SVC(kernel=kerneltype).fit(xtr, ytr)
    w = model.weights()
    if weightfilt:
        opts.append(f"weight filt = {weightfilt}")
        order = torch.argsort(w.abs(), descending=True, stable=True)
        keep = int(round(xtr.shape[1] * weightfilt))
        cols = order[:keep]
        model = SVC(kernel=kerneltype).fit(xtr[:, cols], ytr)
    else:
        opts.append("no weight filt")
    dtr = model.decision_function(xtr[:, cols])
    dte = model.decision_function(xte[:, cols])
    pred = (dte > 0).long()
    rep = M.binary_report(pred, yte)
    return SVMResult(opts, model, w, cols, dtr, dte, pred, M.roc_curve(dte, yte), M.auc(dte, yte), rep.precision, rep.recall,
                     rep.error)


def weight_quantile_tails(weights: torch.Tensor, gene_ids: Sequence[str],
                          fractions: Sequence[float] = (0.25, 0.10, 0.05, 0.01, 0.001)) -> Dict[str, Dict[float, List[str]]]:
    """Genes in the upper / lower ``fractions`` tails of the SVM weight distribution
    (``quantile(weights, 1 - f)`` / ``quantile(weights, f)``; BuieRProj/KM_WHATEVER.Rmd:315-362, which wrote
    High4/10/20/100/1000.csv and low*.csv for f = 25, 10, 5, 1, 0.1 %)."""
    w = weights.double().cpu()
    out = {"high": {}, "low": {}}
    for f in fractions:
        hi = torch.quantile(w, 1 - f)
        lo = torch.quantile(w, f)
        out["high"][f] = [gene_ids[i] for i in torch.nonzero(w > hi).flatten().tolist()]
        out["low"][f] = [gene_ids[i] for i in torch.nonzero(w < lo).flatten().tolist()]
    return out


def grid_search(xtr: torch.Tensor, ytr: torch.Tensor, costs: Sequence[float], gammas: Sequence[float], nfolds: int = 5,
                seed: int = 0) -> Dict[str, torch.Tensor]:
    """RBF-SVM (C, gamma) grid with k-fold CV error; every (C, fold) of one gamma is a CTA of one launch."""
    from ..models.lasso import make_foldid
    n, p = xtr.shape
    dev = xtr.device
    x = xtr.float()
    mu, sd = x.mean(0), x.std(0, unbiased=True)
    xs = ((x - mu) / torch.where(sd > 0, sd, torch.ones_like(sd))).contiguous()
    foldid = torch.as_tensor(make_foldid(n, nfolds, seed), device=dev)
    ypm = 2 * ytr.float() - 1
    err = torch.zeros(len(costs), len(gammas))
    for gi, gamma in enumerate(gammas):
        K = G.rbf_kernel(xs, None, gamma).contiguous()
        cb = torch.stack([torch.where(foldid == f, 0.0, float(c)) for c in costs for f in range(nfolds)]).to(dev)
        alpha, rho, _ = solve_dual(K, ypm, cb)
        coef = alpha.float() * ypm[None, :]
        dec = coef @ K.float().t() - rho.float()[:, None]                     # [C*folds, n]
        dec = dec.reshape(len(costs), nfolds, n)
        held = dec[:, foldid, torch.arange(n, device=dev)]                    # each sample from the fold that left it out
        err[:, gi] = ((held > 0).float() != ytr.float()[None, :]).float().mean(1).cpu()
    best = int(torch.argmin(err))
    return {"cv_error": err, "best_cost": costs[best // len(gammas)], "best_gamma": gammas[best % len(gammas)]}
