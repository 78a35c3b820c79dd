"""Replicate batching (K13): weight rows for bootstrap / subsample / k-fold replicates.

On CUDA the rows come from ``cml_replicate_weights`` (csrc/replicate.cu, counter-based RNG, one warp per
replicate) so that a batch of thousands of replicates never touches the host; on CPU a NumPy generator
produces rows with the same distribution (not the same stream).  Reference call sites that this
replaces are sequential loops: LOOCV inside cv.glmnet (cml_targetaml_seanalysis.Rmd:98-101), repeated CV
(VikasP/AML.R:241-245)."""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _native

MODES = {"bootstrap": 0, "subsample": 1, "kfold": 2}


def replicate_weights(B: int, n: int, *, mode: str = "bootstrap", k: int = 0, seed: int = 0, device="cpu",
                      n_pad: int | None = None, rep0: int = 0) -> torch.Tensor:
    """[B, n_pad] fp32 weights.  ``mode='kfold'``: ``k`` folds, replicate b leaves out fold ``b % k`` of
    repeat ``b // k``.  ``rep0`` offsets the replicate counter so successive batches continue one stream."""
    n_pad = n if n_pad is None else n_pad
    dev = torch.device(device)
    if dev.type == "cuda" or _native.emulating():
        w = torch.empty(B, n_pad, dtype=torch.float32, device=dev)
        rc = _native.cuda().cml_replicate_weights(_native.ptr(w), B, n, n_pad, ctypes.c_ulonglong(seed & (2**64 - 1)),
                                                  MODES[mode], int(k), int(rep0), _native.stream_ptr())
        _native.check(rc, "cml_replicate_weights")
        return w
    rng = np.random.default_rng([seed, rep0])
    w = np.zeros((B, n_pad), np.float32)
    if mode == "bootstrap":
        idx = rng.integers(0, n, size=(B, n))
        np.add.at(w, (np.arange(B)[:, None], idx), 1.0)
    elif mode == "subsample":
        for b in range(B):
            w[b, rng.choice(n, k, replace=False)] = 1.0
    elif mode == "kfold":
        w[:, :n] = 1.0
        nrep = (rep0 + B + k - 1) // k
        folds = [np.arange(n)[np.random.default_rng([seed, r]).permutation(n)] % k for r in range(nrep)]
        for b in range(B):
            r, f = divmod(b + rep0, k)
            w[b, :n][folds[r] == f] = 0.0
    else:
        raise ValueError(mode)
    return torch.from_numpy(w)
