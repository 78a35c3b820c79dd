"""Batched logistic lasso / elastic-net path solver (glmnet semantics).

CUDA tensors are served by ``cml_lognet_path`` (csrc/lasso_path.cu, one warp per replicate, fp32);
CPU tensors by ``cml_host_lognet_path`` (csrc/host/lognet_cpu.cpp, float64), which is also the oracle
the GPU tests compare against.  Replaces ``glmnet()`` as called at
composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:94-95,110 and ``cv.glmnet()`` at :98-101.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from .. import _native


@dataclass
class PathBatch:
    """Compressed coefficient paths of B replicates (glmnet's ``ca / ia / nin`` form)."""

    active: torch.Tensor      # [B, nx] int32   feature index by order of entry
    coef: torch.Tensor        # [B, L, nx]      coefficient of active[b, j] at lambda l (valid for j < kin[b, l])
    kin: torch.Tensor         # [B, L] int32
    a0: torch.Tensor          # [B, L]
    dev_ratio: torch.Tensor   # [B, L]
    eta: Optional[torch.Tensor]  # [B, L, n] linear predictor of every sample (incl. weighted-out ones)
    nulldev: torch.Tensor     # [B]
    npasses: torch.Tensor     # [B] int32
    lmu: torch.Tensor         # [B] int32  number of lambdas actually fitted
    jerr: torch.Tensor        # [B] int32
    lam: torch.Tensor         # [L]
    p: int

    @property
    def B(self) -> int:
        return self.coef.shape[0]

    def dense(self, b: int = 0) -> torch.Tensor:
        """[L, p] dense coefficients of replicate ``b``."""
        L, nx = self.coef.shape[1], self.coef.shape[2]
        out = torch.zeros(L, self.p, dtype=self.coef.dtype, device=self.coef.device)
        kin = self.kin[b].long()
        kmax = int(kin.max().item()) if L else 0
        if kmax == 0:
            return out
        idx = self.active[b, :kmax].long()
        mask = torch.arange(kmax, device=self.coef.device)[None, :] < kin[:, None]
        vals = torch.where(mask, self.coef[b, :, :kmax], torch.zeros((), dtype=self.coef.dtype, device=self.coef.device))
        out[:, idx] = vals
        return out

    def df(self) -> torch.Tensor:
        """[B, L] number of non-zero coefficients."""
        nx = self.coef.shape[2]
        mask = torch.arange(nx, device=self.coef.device)[None, None, :] < self.kin[:, :, None]
        return ((self.coef != 0) & mask).sum(-1)

    def coef_at(self, b: int, l: int) -> torch.Tensor:
        return self.dense(b)[l]


def pack_design(x: torch.Tensor) -> tuple[torch.Tensor, int]:
    """samples x genes -> the kernel's feature-major, zero-padded layout ``[p, n_pad]`` (fp32)."""
    n, p = x.shape
    n_pad = _n_pad(n)
    xt = torch.zeros(p, n_pad, dtype=torch.float32, device=x.device)
    xt[:, :n] = x.t().to(torch.float32)
    return xt, n_pad


def _n_pad(n: int) -> int:
    for s in (32, 64, 96, 128, 160, 192, 256, 384, 512, 768, 1024):
        if n <= s:
            return s
    raise ValueError(f"the warp-per-replicate path solver handles n <= 1024 samples (got {n}); "
                     "use consensusml_b200.models.lasso_large for sample-sharded cohorts")


def default_lambda_grid() -> np.ndarray:
    """The reference's grid ``10^seq(10, -2, length=100)`` (cml_targetaml_seanalysis.Rmd:92)."""
    return 10.0 ** np.linspace(10.0, -2.0, 100)


def lognet_path(x: torch.Tensor, y: torch.Tensor, weights: torch.Tensor, lam, *, alpha: float = 1.0,
                intercept: bool = False, modified_newton: bool = False, thresh: float = 1e-7,
                maxit: int = 100000, nx: Optional[int] = None, feature_mask: Optional[torch.Tensor] = None,
                penalty_factor: Optional[torch.Tensor] = None, want_eta: bool = True,
                packed: Optional[tuple] = None, freq_out: Optional[torch.Tensor] = None,
                padded_weights: Optional[torch.Tensor] = None, want_coef: bool = True) -> PathBatch:
    """Fit B replicate paths that share the design matrix.

    x : [n, p] samples x genes;  y : [n] in {0,1} (or [B, n]);  weights : [B, n] (0 leaves a sample out);
    lam : descending lambda grid;  feature_mask : [p] or [B, p] bool (False = gene excluded);
    freq_out : optional [L, p] fp32 accumulator of selection counts (stability selection).
    """
    if weights.dim() == 1:
        weights = weights[None, :]
    B, n = weights.shape
    p = x.shape[1]
    lam_np = np.asarray(lam, dtype=np.float64)
    if np.any(np.diff(lam_np) > 0):
        raise ValueError("lambda grid must be descending")
    L = len(lam_np)
    nx = int(min(p, nx if nx is not None else 256))
    if penalty_factor is not None:
        pf = penalty_factor.to(torch.float64).clamp_min(0)
        penalty_factor = pf * p / pf.sum()
    if freq_out is not None and (tuple(freq_out.shape) != (L, p) or freq_out.dtype != torch.float32):
        raise ValueError("freq_out must be a float32 [L, p] tensor")
    if _native.on_device(x):
        if x.is_cuda:
            _latency_kernel_canary(x.device)
        return _path_cuda(x, y, weights, lam_np, alpha, intercept, modified_newton, thresh, maxit, nx,
                          feature_mask, penalty_factor, want_eta, packed, freq_out, padded_weights, want_coef)
    out = _path_host(x, y, weights, lam_np, alpha, intercept, modified_newton, thresh, maxit, nx,
                     feature_mask, penalty_factor, want_eta)
    if freq_out is not None:
        freq_out += (out._dense_all != 0).sum(0).to(freq_out.dtype)
    return out


_CANARY_DONE: dict = {}


def _latency_kernel_canary(device) -> None:
    """Once per process and device: a small batch through BOTH path kernels.  The one-CTA-per-fit kernel (latency mode, chosen
    for batches of at most one fit per SM) was developed on the host execution model, where it is bit-identical to the
    one-warp kernel; this makes the same comparison on the actual device before the pipeline relies on it.  A mismatch
    switches latency mode off for the process (``CML_LASSO_TEAM=0``) and says so - results then come from the kernel with
    the hardware record.  An explicit ``CML_LASSO_TEAM`` setting is respected and skips the check."""
    import os
    import warnings
    key = str(device)
    if key in _CANARY_DONE or "CML_LASSO_TEAM" in os.environ:
        return
    _CANARY_DONE[key] = True
    g = torch.Generator().manual_seed(11)
    n, p, B = 60, 200, 4
    x = torch.randn(n, p, generator=g)
    bt = torch.zeros(p)
    bt[:5] = torch.tensor([1.0, -1.0, 0.8, -0.6, 0.5])
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ bt)).float()
    w = torch.ones(B, n)
    for b in range(1, B):
        w[b, torch.randperm(n, generator=g)[:6]] = 0
    lam = 10.0 ** np.linspace(-0.4, -1.8, 25)
    outs = []
    try:
        for mode in ("0", "1"):
            os.environ["CML_LASSO_TEAM"] = mode
            pb = _path_cuda(x.to(device), y.to(device), w.to(device), lam, 1.0, False, False, 1e-7, 100000, min(p, 256), None, None, False, None)
            outs.append((pb.npasses.cpu(), pb.kin.cpu(), torch.stack([pb.dense(b).cpu() for b in range(B)]), pb.jerr.cpu()))
    finally:
        os.environ.pop("CML_LASSO_TEAM", None)
    same = all(torch.equal(a, b) for a, b in zip(*outs)) and int(outs[0][3].abs().max()) == 0
    if not same:
        os.environ["CML_LASSO_TEAM"] = "0"
        warnings.warn("consensusml_b200: the latency-mode lasso kernel disagreed with the one-warp kernel on this device; "
                      "latency mode is switched off for this process (CML_LASSO_TEAM=0)")
    _CANARY_DONE[key] = "ok" if same else "disabled"


def _path_cuda(x, y, weights, lam_np, alpha, intercept, kopt, thresh, maxit, nx, fmask, vp, want_eta, packed,
               freq_out=None, padded_weights=None, want_coef=True):
    lib = _native.cuda()
    dev = x.device
    B, n = weights.shape
    p = x.shape[1]
    L = len(lam_np)
    if packed is not None:
        xt, n_pad = packed
    else:
        xt, n_pad = pack_design(x)
    f32 = dict(dtype=torch.float32, device=dev)
    i32 = dict(dtype=torch.int32, device=dev)
    if padded_weights is not None and tuple(padded_weights.shape) == (B, n_pad) and padded_weights.dtype == torch.float32:
        wp = padded_weights.contiguous()          # already in the kernel's layout: no copy
    else:
        wp = torch.zeros(B, n_pad, **f32)
        wp[:, :n] = weights.to(**f32)
    if y.dim() == 1:
        yp = torch.zeros(n_pad, **f32)
        yp[:n] = y.to(**f32)
        y_stride = 0
    else:
        yp = torch.zeros(B, n_pad, **f32)
        yp[:, :n] = y.to(**f32)
        y_stride = n_pad
    ju = None
    ju_stride = 0
    if fmask is not None:
        ju = fmask.to(device=dev, dtype=torch.uint8).contiguous()
        ju_stride = p if ju.dim() == 2 else 0
    vpt = vp.to(**f32).contiguous() if vp is not None else None
    lam_t = torch.as_tensor(lam_np, **f32)
    lib.cml_lognet_scratch_bytes.restype = ctypes.c_size_t
    sbytes = lib.cml_lognet_scratch_bytes(B, n_pad, p)
    scratch = torch.empty(sbytes, dtype=torch.uint8, device=dev)
    m_out = torch.zeros(B, nx, **i32) if want_coef else None
    ca = torch.empty(B, L, nx, **f32) if want_coef else None       # entries >= kin[b, l] are never read
    kin = torch.zeros(B, L, **i32)
    a0 = torch.zeros(B, L, **f32)
    devr = torch.zeros(B, L, **f32)
    eta = torch.zeros(B, L, n_pad, **f32) if want_eta else None
    nulldev = torch.zeros(B, **f32)
    npasses = torch.zeros(B, **i32)
    lmu = torch.zeros(B, **i32)
    jerr = torch.zeros(B, **i32)
    P = _native.ptr
    rc = lib.cml_lognet_path(
        B, n, n_pad, p, L, nx, P(xt), P(yp), ctypes.c_longlong(y_stride), P(wp), P(ju),
        ctypes.c_longlong(ju_stride), P(vpt), P(lam_t), ctypes.c_float(alpha), ctypes.c_float(thresh),
        int(bool(intercept)), int(bool(kopt)), int(maxit), P(scratch), ctypes.c_size_t(sbytes), P(m_out), P(ca),
        P(kin), P(a0), P(devr), P(eta), P(freq_out), P(nulldev), P(npasses), P(lmu), P(jerr), _native.stream_ptr())
    _native.check(rc, "cml_lognet_path")
    if not want_coef:
        m_out = torch.zeros(B, 0, **i32)
        ca = torch.zeros(B, L, 0, **f32)
        kin = torch.zeros_like(kin)
    return PathBatch(m_out, ca, kin, a0, devr, eta[:, :, :n] if eta is not None else None, nulldev, npasses, lmu,
                     jerr, lam_t, p)


def _path_host(x, y, weights, lam_np, alpha, intercept, kopt, thresh, maxit, nx, fmask, vp, want_eta):
    lib = _native.host()
    B, n = weights.shape
    p = x.shape[1]
    L = len(lam_np)
    xt = np.ascontiguousarray(x.detach().cpu().numpy().T.astype(np.float64))
    w = np.ascontiguousarray(weights.detach().cpu().numpy().astype(np.float64))
    lam_np = np.ascontiguousarray(lam_np)
    vp_np = np.ascontiguousarray(vp.cpu().numpy().astype(np.float64)) if vp is not None else None
    ys = y.detach().cpu().numpy().astype(np.float64)
    per_rep_y = ys.ndim == 2
    ju_np = None
    if fmask is not None:
        ju_np = fmask.detach().cpu().numpy().astype(np.int8)
        if ju_np.ndim == 1:
            ju_np = np.broadcast_to(ju_np, (B, p))
        ju_np = np.ascontiguousarray(ju_np)
    beta = np.zeros((B, L, p))
    a0 = np.zeros((B, L))
    dev = np.zeros((B, L))
    df = np.zeros((B, L), np.int32)
    nulldev = np.zeros(B)
    npasses = np.zeros(B, np.int32)
    lmu = np.zeros(B, np.int32)
    jerr = np.zeros(B, np.int32)
    P = _native.np_ptr

    def call(b0, b1, yv):
        nb = b1 - b0
        lib.cml_host_lognet_path(
            nb, n, p, L, P(xt), P(np.ascontiguousarray(yv)), P(w[b0:b1]), P(ju_np[b0:b1]) if ju_np is not None else None,
            P(vp_np), P(lam_np), ctypes.c_double(alpha), ctypes.c_double(thresh), int(bool(intercept)), int(bool(kopt)),
            int(maxit), int(p), P(beta[b0:b1]), P(a0[b0:b1]), P(dev[b0:b1]), P(df[b0:b1]), P(nulldev[b0:b1]),
            P(npasses[b0:b1]), P(lmu[b0:b1]), P(jerr[b0:b1]))

    if per_rep_y:
        for b in range(B):
            call(b, b + 1, ys[b])
    else:
        call(0, B, ys)
    # compress into the same form the CUDA path returns
    nz = (beta != 0).any(axis=1)                        # [B, p]
    kmax = max(1, int(nz.sum(1).max()))
    nxo = max(1, min(max(nx, kmax), p))
    active = np.zeros((B, nxo), np.int32)
    coef = np.zeros((B, L, nxo))
    kin = np.zeros((B, L), np.int32)
    for b in range(B):
        idx = np.flatnonzero(nz[b])
        # order of entry = first lambda at which the coefficient is non-zero (ties by index)
        first = np.array([np.argmax(beta[b, :, j] != 0) for j in idx], dtype=np.int64)
        order = np.lexsort((idx, first))
        idx = idx[order]
        active[b, :len(idx)] = idx
        coef[b, :, :len(idx)] = beta[b][:, idx]
        ever = np.maximum.accumulate((beta[b][:, idx] != 0), axis=0) if len(idx) else np.zeros((L, 0), bool)
        kin[b] = ever.sum(1)
    eta = None
    if want_eta:
        xs = x.detach().cpu().numpy().astype(np.float64)
        eta = np.zeros((B, L, n))
        for b in range(B):
            k = int(kin[b].max())
            if k:
                eta[b] = coef[b, :, :k] @ xs[:, active[b, :k]].T
        eta += a0[:, :, None]
    t = lambda a, dt=torch.float64: torch.as_tensor(a, dtype=dt)
    out = PathBatch(t(active, torch.int32), t(coef), t(kin, torch.int32), t(a0), t(dev), t(eta) if eta is not None else None,
                    t(nulldev), t(npasses, torch.int32), t(lmu, torch.int32), t(jerr, torch.int32), t(lam_np), p)
    out._dense_all = torch.as_tensor(beta)
    return out
