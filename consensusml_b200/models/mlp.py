"""Small dense networks trained over a hyper-parameter grid in one batch (K17).

The reference advertises "Neural Networks with keras" (README.md:25) but ships no such code; BASELINE.json asks for a
"small Keras-style MLP" trained across its hyper-parameter grid.  All G configurations (hidden width, learning rate,
weight decay, seed) are trained together: their first layers are concatenated along the output axis, so the forward
p -> sum(h) product and the matching weight-gradient are two tcgen05 GEMMs per step on CUDA
(``ops.gemm.gemm_tn``, bf16 operands / fp32 accumulate, fp32 master weights).  Everything else of an epoch is two kernels
of csrc/mlp.cu: ``mlp_tail_kernel`` (bias + ReLU + per-configuration output unit + sigmoid + BCE and its gradient, writing
dH^T in bf16 as GEMM 2's operand) and ``mlp_adam_kernel`` (Adam for all four parameter groups, small-gradient reductions,
bf16 refresh of W1) - four launches per epoch; the standardised inputs are converted to bf16 once.  Sample-sharded
training all-reduces the gradients once per step.  CPU tensors run the same maths in fp32 torch (also the oracle)."""
from __future__ import annotations

from dataclasses import dataclass
from itertools import product
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

import ctypes

from .. import _native
from ..ops import gemm as G
from ..parallel.dist import all_reduce_

_I, _LL, _F, _P = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p
_native.register_signature("cml_mlp_tail", [_P, _LL, _I, _I, _I, _P, _P, _P, _P, _P, _F, _P, _LL, _P, _P, _P, _P, _P])
_native.register_signature("cml_mlp_adam", [_P, _P, _LL, _P, _P, _P, _P, _LL, _P, _P, _P, _I, _I, _I, _I, _F, _F, _F, _I, _P])


@dataclass
class MLPGridResult:
    configs: List[Dict]
    train_loss: torch.Tensor       # [G]
    val_loss: Optional[torch.Tensor]
    val_auc: Optional[torch.Tensor]
    w1: torch.Tensor               # [H_total, p]
    b1: torch.Tensor
    w2: torch.Tensor               # [H_total]
    b2: torch.Tensor               # [G]
    offsets: List[int]
    mean: torch.Tensor
    scale: torch.Tensor

    def predict_proba(self, x: torch.Tensor, config: int) -> torch.Tensor:
        xs = ((x.float() - self.mean) / self.scale)
        lo, hi = self.offsets[config], self.offsets[config + 1]
        h = torch.relu(xs @ self.w1[lo:hi].t() + self.b1[lo:hi])
        return torch.sigmoid(h @ self.w2[lo:hi] + self.b2[config])

    def input_importance(self, config: int) -> torch.Tensor:
        """|w2|-weighted column norms of the first layer: a per-gene importance for the consensus table."""
        lo, hi = self.offsets[config], self.offsets[config + 1]
        return (self.w1[lo:hi].abs() * self.w2[lo:hi].abs()[:, None]).sum(0)


def make_grid(hidden=(16, 32, 64), lr=(1e-3, 3e-3), weight_decay=(0.0, 1e-3), seeds=(0,)) -> List[Dict]:
    return [dict(hidden=h, lr=l, weight_decay=w, seed=s) for h, l, w, s in product(hidden, lr, weight_decay, seeds)]


def _gemm_nt(a_bf16: torch.Tensor, b_bf16: torch.Tensor, n_out: int) -> torch.Tensor:
    """A . B^T for bf16 [M, K] x [N, K] -> fp32 [M, >= n_out] on the tcgen05 kernel (GPU, or the host execution model's
    rendering of it); toy shapes below one k-block take the plain product of the same bf16 values."""
    if _native.on_device(a_bf16) and a_bf16.shape[1] >= G.MIN_K and a_bf16.shape[0] >= 32 and n_out >= 32:
        return G.gemm_tn(a_bf16, [b_bf16], n_out)
    return a_bf16.float() @ b_bf16.float().t()


def _bf16_rows(t: torch.Tensor) -> torch.Tensor:
    """fp32 [r, c] -> bf16 [r, pad8(c)] zero padded (16-byte rows for TMA)."""
    r, c = t.shape
    out = torch.zeros(r, (c + 7) // 8 * 8, dtype=torch.bfloat16, device=t.device)
    out[:, :c] = t.to(torch.bfloat16)
    return out


def _train_epochs_kernels(xs, yf, w1, b1, w2, b2, m, v, seg, offsets, configs, epochs, n_total, sharded, beta1, beta2, eps):
    """The epoch loop on the kernel path: GEMM 1 -> mlp_tail -> GEMM 2 -> mlp_adam (csrc/mlp.cu)."""
    dev = xs.device
    n, p = xs.shape
    H, Gn = w1.shape[0], len(configs)
    lib = _native.cuda()
    P = _native.ptr
    x_bf = _bf16_rows(xs)                                   # [n, p8]   converted ONCE
    xt_bf = _bf16_rows(xs.t().contiguous())                 # [p, n8]   B operand of GEMM 2
    w1_bf = _bf16_rows(w1)                                  # [H, p8]   refreshed by the Adam kernel
    ldn = xt_bf.shape[1]
    dhT = torch.zeros(H, ldn, dtype=torch.bfloat16, device=dev)     # padding columns stay zero
    hv = torch.empty(n, H, device=dev)
    dl = torch.empty(n, Gn, device=dev)
    loss = torch.empty(n, Gn, device=dev)
    seg_off = torch.tensor(offsets, dtype=torch.int32, device=dev)
    cfg_of = seg.to(torch.int32).contiguous()
    lr = torch.tensor([c["lr"] for c in configs], dtype=torch.float32, device=dev)
    wd = torch.tensor([c.get("weight_decay", 0.0) for c in configs], dtype=torch.float32, device=dev)
    state = [w1, b1, w2, b2] + m + v
    state_ptrs = (ctypes.c_void_p * 12)(*[t.data_ptr() for t in state])
    inv = 1.0 / float(n_total)
    for t in range(1, epochs + 1):
        pre = _gemm_nt(x_bf, w1_bf, H)
        rc = lib.cml_mlp_tail(P(pre), pre.stride(0), n, H, Gn, P(seg_off), P(b1), P(w2), P(b2), P(yf), inv, P(dhT), ldn, P(hv), P(dl), P(loss),
                              None, _native.stream_ptr())
        _native.check(rc, "cml_mlp_tail")
        gw1 = _gemm_nt(dhT, xt_bf, p)
        gsmall = None
        if sharded:
            act = hv > 0
            dseg = dl[:, seg]
            gsmall = torch.cat([(dseg * w2[None, :] * act).sum(0), (dseg * hv).sum(0), dl.sum(0)]).contiguous()
            gw1 = gw1.contiguous()
            all_reduce_(gw1)
            all_reduce_(gsmall)
        rc = lib.cml_mlp_adam(state_ptrs, P(gw1), gw1.stride(0), P(hv), P(dl), P(gsmall), P(w1_bf), w1_bf.shape[1], P(cfg_of), P(lr), P(wd),
                              n, H, Gn, p, beta1, beta2, eps, t, _native.stream_ptr())
        _native.check(rc, "cml_mlp_adam")
    return loss.sum(0)


def train_grid(x: torch.Tensor, y: torch.Tensor, configs: Sequence[Dict], epochs: int = 200, x_val: Optional[torch.Tensor] = None,
               y_val: Optional[torch.Tensor] = None, n_total: Optional[int] = None, sharded: bool = False) -> MLPGridResult:
    """Full-batch Adam on binary cross-entropy for every configuration at once (one hidden ReLU layer).
    ``sharded=True``: ``x`` holds this rank's samples of an ``n_total``-sample cohort and the gradients are all-reduced
    (a collective - every rank must call); the default trains on ``x`` alone, whatever process group exists."""
    from ..ops.metrics import auc
    dev = x.device
    n, p = x.shape
    n_total = n if n_total is None else n_total
    xf = x.float()
    mean = xf.mean(0)
    sd = xf.std(0, unbiased=True)
    scale = torch.where(sd > 0, sd, torch.ones_like(sd))
    xs = ((xf - mean) / scale).contiguous()
    yf = y.float()
    Gn = len(configs)
    hs = [int(c["hidden"]) for c in configs]
    offsets = [0] + list(np.cumsum(hs))
    H = offsets[-1]
    w1 = torch.empty(H, p, device=dev)
    w2 = torch.empty(H, device=dev)
    for gi, c in enumerate(configs):
        gen = torch.Generator(device="cpu").manual_seed(int(c.get("seed", 0)) * 1009 + gi)
        lo, hi = offsets[gi], offsets[gi + 1]
        w1[lo:hi] = (torch.randn(hi - lo, p, generator=gen) * np.sqrt(2.0 / p)).to(dev)
        w2[lo:hi] = (torch.randn(hi - lo, generator=gen) * np.sqrt(1.0 / (hi - lo))).to(dev)
    b1 = torch.zeros(H, device=dev)
    b2 = torch.zeros(Gn, device=dev)
    seg = torch.repeat_interleave(torch.arange(Gn, device=dev), torch.tensor(hs, device=dev))     # hidden unit -> config
    lr_u = torch.tensor([c["lr"] for c in configs], device=dev)[seg]
    wd_u = torch.tensor([c.get("weight_decay", 0.0) for c in configs], device=dev)[seg]
    lr_g = torch.tensor([c["lr"] for c in configs], device=dev)
    params = [w1, b1, w2, b2]
    m = [torch.zeros_like(t) for t in params]
    v = [torch.zeros_like(t) for t in params]
    lrs = [lr_u[:, None], lr_u, lr_u, lr_g]
    wds = [wd_u[:, None], None, wd_u, None]
    b1a, b2a, eps = 0.9, 0.999, 1e-8
    loss_g = torch.zeros(Gn, device=dev)
    if _native.on_device(x):
        loss_g = _train_epochs_kernels(xs, yf, w1, b1, w2, b2, m, v, seg, offsets, configs, epochs, n_total, sharded, b1a, b2a, eps)
    else:
        xst = xs.t().contiguous()
        for t in range(1, epochs + 1):
            pre = xs @ w1.t() + b1                                         # [n, H]   GEMM 1
            h = torch.relu(pre)
            logit = torch.zeros(n, Gn, device=dev).index_add_(1, seg, h * w2[None, :]) + b2[None, :]
            prob = torch.sigmoid(logit)
            dlogit = (prob - yf[:, None]) / n_total                        # [n, G]
            loss_g = -(yf[:, None] * torch.log(prob.clamp_min(1e-12)) + (1 - yf[:, None]) * torch.log((1 - prob).clamp_min(1e-12))).sum(0)
            dh = dlogit[:, seg] * w2[None, :] * (pre > 0)                  # [n, H]
            gw2 = (dlogit[:, seg] * h).sum(0)
            gb2 = dlogit.sum(0)
            gw1 = dh.t().contiguous() @ xst.t()                            # [H, p]   GEMM 2
            gb1 = dh.sum(0)
            grads = [gw1, gb1, gw2, gb2]
            for gr in (grads if sharded else ()):
                all_reduce_(gr)                                            # sample-sharded: one reduction per tensor
            for i, (pt, gr) in enumerate(zip(params, grads)):
                if wds[i] is not None:
                    gr = gr + wds[i] * pt
                m[i].mul_(b1a).add_(gr, alpha=1 - b1a)
                v[i].mul_(b2a).addcmul_(gr, gr, value=1 - b2a)
                mh = m[i] / (1 - b1a ** t)
                vh = v[i] / (1 - b2a ** t)
                pt.sub_(lrs[i] * mh / (vh.sqrt() + eps))
    if sharded:
        all_reduce_(loss_g)
    res = MLPGridResult(list(configs), loss_g / n_total, None, None, w1, b1, w2, b2, [int(o) for o in offsets], mean, scale)
    if x_val is not None:
        yv = y_val.float()
        vl, va = [], []
        for gi in range(Gn):
            pr = res.predict_proba(x_val, gi).clamp(1e-12, 1 - 1e-12)
            vl.append(float(-(yv * torch.log(pr) + (1 - yv) * torch.log(1 - pr)).mean()))
            va.append(auc(pr, yv))
        res.val_loss, res.val_auc = torch.tensor(vl), torch.tensor(va)
    return res
