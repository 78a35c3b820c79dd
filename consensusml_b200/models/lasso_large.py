"""Sample-sharded L1 logistic regression for cohorts too large for the warp-per-replicate solver.

The reference's cohorts have n <= 474 samples; the north-star asks for 1M samples x 20k genes across 8 GPUs.
At that size the design matrix is streamed, not cached: every rank holds a contiguous block of samples (bf16,
row-major) and the genes are partitioned among the ranks for the update.  One proximal-gradient (FISTA)
iteration is

    r      = w (y - sigmoid(X v))                 xb_residual kernel        (local rows)
    g_own  = sum_ranks X_r^T r_r  [own genes]     xtr_reduce_scatter kernel (FUSED: peer red.add + signals)
    beta   = soft(v + s g/n, s lambda)            prox_allgather kernel     (FUSED: peer stores of the new slice)

so the two collectives of the iteration (reduce-scatter of the gradient, all-gather of the coefficients) are
issued from inside the kernels over NVLink.  ``backend="unfused"`` runs the same iteration with torch ops and
``torch.distributed`` collectives (NCCL or gloo) - the comparator and the CPU path.

glmnet parity note: this solver minimises the same objective ((1/n) sum w_i logloss_i + lambda |beta|_1, no
intercept by default) but by FISTA instead of cyclic coordinate descent, so it agrees with the path solver to
optimisation tolerance, not update for update."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from .. import _native
from ..parallel import fused as F
from ..parallel import symm
from ..parallel.dist import all_reduce_, equal_shard


@dataclass
class LargeFit:
    beta: torch.Tensor
    iterations: int
    objective: float
    step: float
    history: list


def lipschitz_estimate(x: torch.Tensor, n_total: int, iters: int = 12) -> float:
    """Largest eigenvalue of X^T X / (4 n) by power iteration (set-up only; partial products all-reduced)."""
    p = x.shape[1]
    g = torch.Generator(device="cpu").manual_seed(0)
    v = torch.randn(p, generator=g).to(x.device)
    v /= v.norm()
    lam = 1.0
    xf = x if x.dtype == torch.float32 else None
    for _ in range(iters):
        u = (x.float() @ v) if xf is None else xf @ v
        w = (x.float().t() @ u) if xf is None else xf.t() @ u
        all_reduce_(w)
        lam = float(w.norm())
        v = w / (lam + 1e-30)
    return lam / (4.0 * n_total)


def fit(x_local: torch.Tensor, y_local: torch.Tensor, lam: float, *, n_total: Optional[int] = None, max_iter: int = 200,
        tol: float = 1e-5, backend: str = "auto", weights: Optional[torch.Tensor] = None, beta0: Optional[torch.Tensor] = None,
        step: Optional[float] = None, check_every: int = 10, alpha: float = 1.0, deterministic: bool = False) -> LargeFit:
    """L1 (``alpha = 1``) or elastic-net logistic regression, glmnet's penalty ``lam (alpha |b|_1 + (1 - alpha)/2 |b|^2)``, on this
    rank's rows (``x_local`` [n_r, p]); every rank returns the full ``beta``.  ``deterministic=True`` runs the gradient
    reduce-scatter in its fixed-order form (no float atomics: bit-identical fits from run to run)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    n_r, p = x_local.shape
    dev = x_local.device
    if n_total is None:
        t = torch.tensor([n_r], dtype=torch.float64, device=dev)
        all_reduce_(t)
        n_total = int(t.item())
    if backend == "auto":
        backend = "fused" if _native.on_device(x_local) else "unfused"
    if step is None:
        step = 1.0 / max(lipschitz_estimate(x_local, n_total), 1e-12)
    if backend == "unfused":
        return _fit_unfused(x_local, y_local, lam, n_total, max_iter, tol, weights, beta0, step, check_every, alpha)

    lb = _native.cuda()
    P = _native.ptr
    st = _native.stream_ptr
    xb = x_local if x_local.dtype == torch.bfloat16 else F.to_bf16_padded(x_local)
    ldx = xb.shape[1]
    rs = F.XtrReduceScatter(p, 1, dev, deterministic=deterministic, n_rows=n_r)
    rank = rs.rank
    beta = symm.alloc((ldx,), torch.float32, dev)
    v = symm.alloc((ldx,), torch.float32, dev)
    if beta0 is not None:
        beta.local[:p] = beta0.float()
        v.local[:p] = beta0.float()
    y = y_local.float().contiguous()
    w = weights.float().contiguous() if weights is not None else None
    R = torch.empty(n_r, dtype=torch.float32, device=dev)
    loss = torch.zeros(1, dtype=torch.float64, device=dev)
    done = torch.zeros(4, dtype=torch.int32, device=dev)
    dmax = torch.zeros(1, dtype=torch.float32, device=dev)
    _native.synchronize()
    beta.barrier()
    bptr, vptr, sptr = F._ptr_array(beta.peer_ptrs), F._ptr_array(v.peer_ptrs), F._ptr_array(rs.grad.signal_ptrs)
    my_pad = ctypes.c_void_p(rs.grad.signal_ptrs[rank])
    tk = 1.0
    history = []
    it = 0
    obj = float("nan")
    for it in range(1, max_iter + 1):
        if it > 1:      # the coefficient slices of every owner have landed in this rank's copy
            _native.check(lb.cml_wait_signals(my_pad, F.SIG_BETA, world, it - 1, st()), "cml_wait_signals")
        check = (it % check_every == 0) or it == max_iter
        if check:
            loss.zero_()
            dmax.zero_()
        _native.check(lb.cml_xb_residual(P(xb), ldx, n_r, p, P(v.local), 0.0, P(y), P(w), P(R), None, P(loss), st()),
                      "cml_xb_residual")
        rs.launch(xb, R)
        rs.wait()
        tk1 = (1.0 + np.sqrt(1.0 + 4.0 * tk * tk)) / 2.0
        mom = (tk - 1.0) / tk1
        tk = tk1
        if alpha >= 1.0:
            _native.check(lb.cml_prox_allgather(P(rs.grad.local), rs.genes_owned, rs.gene0, None, step, lam, 1.0 / n_total, mom,
                                                bptr, vptr, sptr, rank, world, P(beta.local), P(done), it, F.SIG_BETA, P(dmax),
                                                st()), "cml_prox_allgather")
        else:
            _native.check(lb.cml_prox_allgather_enet(P(rs.grad.local), rs.genes_owned, rs.gene0, None, step, lam * alpha,
                                                     lam * (1.0 - alpha), 1.0 / n_total, mom, bptr, vptr, sptr, rank, world,
                                                     P(beta.local), P(done), it, F.SIG_BETA, P(dmax), st()), "cml_prox_allgather_enet")
        if check:
            # convergence test on the largest coefficient change of this iteration (max over ranks); the loss printed
            # is that of the extrapolated point v, enough to watch progress
            stats = torch.stack([dmax[0].double(), loss[0]])
            d = stats[:1].clone()
            s = stats[1:].clone()
            all_reduce_(d, dist.ReduceOp.MAX)
            all_reduce_(s)
            _native.check(lb.cml_wait_signals(my_pad, F.SIG_BETA, world, it, st()), "cml_wait_signals")
            bl = beta.local[:p]
            obj = float(s.item()) / n_total + lam * (alpha * float(bl.abs().sum()) + 0.5 * (1.0 - alpha) * float((bl * bl).sum()))
            history.append((it, obj, float(d.item())))
            if float(d.item()) < tol:
                break
    _native.check(lb.cml_wait_signals(my_pad, F.SIG_BETA, world, it, st()), "cml_wait_signals")
    _native.synchronize()
    beta.barrier()
    out = beta.local[:p].clone()
    _native.synchronize()
    beta.barrier()
    return LargeFit(out, it, obj, step, history)


def _fit_unfused(x, y, lam, n_total, max_iter, tol, weights, beta0, step, check_every, alpha: float = 1.0) -> LargeFit:
    p = x.shape[1]
    dev = x.device
    xf = x.float()
    beta = torch.zeros(p, dtype=torch.float32, device=dev) if beta0 is None else beta0.float().clone()
    v = beta.clone()
    w = weights.float() if weights is not None else torch.ones_like(y, dtype=torch.float32)
    yf = y.float()
    tk = 1.0
    history = []
    obj = float("nan")
    it = 0
    for it in range(1, max_iter + 1):
        eta = xf @ v
        r = w * (yf - torch.sigmoid(eta))
        g = xf.t() @ r
        all_reduce_(g)
        z = v + step * g / n_total
        thr = step * lam * alpha
        bnew = torch.sign(z) * (z.abs() - thr).clamp_min(0) / (1.0 + step * lam * (1.0 - alpha))
        tk1 = (1.0 + np.sqrt(1.0 + 4.0 * tk * tk)) / 2.0
        mom = (tk - 1.0) / tk1
        tk = tk1
        d = float((bnew - beta).abs().max())
        v = bnew + mom * (bnew - beta)
        beta = bnew
        if it % check_every == 0 or it == max_iter:
            sp = torch.where(eta > 0, eta + torch.log1p(torch.exp(-eta)), torch.log1p(torch.exp(eta)))
            s = (w * (sp - yf * eta)).double().sum().reshape(1)
            all_reduce_(s)
            obj = float(s.item()) / n_total + lam * (alpha * float(beta.abs().sum()) + 0.5 * (1.0 - alpha) * float((beta * beta).sum()))
            history.append((it, obj, d))
            if d < tol:
                break
    return LargeFit(beta, it, obj, step, history)


def lambda_max(x_local: torch.Tensor, y_local: torch.Tensor, n_total: int, weights: Optional[torch.Tensor] = None) -> float:
    """Smallest lambda with an all-zero solution (no intercept): ``max_j |x_j^T w (y - 1/2)| / n`` (one all-reduce)."""
    w = weights.float() if weights is not None else torch.ones_like(y_local, dtype=torch.float32)
    g = x_local.float().t() @ (w * (y_local.float() - 0.5))
    all_reduce_(g)
    return float(g.abs().max()) / n_total


def fit_path(x_local: torch.Tensor, y_local: torch.Tensor, lams=None, *, n_lambda: int = 20, lambda_min_ratio: float = 0.05,
             n_total: Optional[int] = None, max_iter: int = 300, tol: float = 1e-5, backend: str = "auto",
             weights: Optional[torch.Tensor] = None, alpha: float = 1.0):
    """Warm-started path of the sharded solver over a descending lambda grid (default: geometric from ``lambda_max`` down to
    ``lambda_min_ratio * lambda_max``, glmnet's construction).  The Lipschitz step is estimated once.  Returns
    ``(lams, [LargeFit per lambda])``; every rank holds every coefficient vector."""
    if n_total is None:
        t = torch.tensor([x_local.shape[0]], dtype=torch.float64, device=x_local.device)
        all_reduce_(t)
        n_total = int(t.item())
    if lams is None:
        lmax = lambda_max(x_local, y_local, n_total, weights) / max(alpha, 1e-3)
        lams = np.geomspace(lmax, lmax * lambda_min_ratio, n_lambda)
    lams = np.asarray(lams, dtype=np.float64)
    if np.any(np.diff(lams) > 0):
        raise ValueError("lambda grid must be descending")
    step = 1.0 / max(lipschitz_estimate(x_local, n_total), 1e-12)
    fits, beta = [], None
    for lam in lams:
        f = fit(x_local, y_local, float(lam), n_total=n_total, max_iter=max_iter, tol=tol, backend=backend, weights=weights,
                beta0=beta, step=step, alpha=alpha)
        beta = f.beta
        fits.append(f)
    return lams, fits


# ------------------------------------------------------------------------------------------------ replicate-batched form
_I, _LL, _Fc, _Pc, _U = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p, ctypes.c_uint
_native.register_signature("cml_batched_residual", [_Pc, _LL, _I, _I, _Pc, _Pc, _LL, _Pc, _LL, _Pc, _Pc])
_native.register_signature("cml_batched_prox_allgather", [_Pc, _LL, _I, _I, _I, _Pc, _Fc, _Fc, _Fc, _Fc, _Pc, _Pc, _LL, _Pc, _LL, _Pc, _I, _I,
                                                            _Pc, _U, _I, _Pc, _Pc])
SIG_VHL = 48                    # signal-pad words 48..55 (fused.py: 0..23, transpose: 24..31, gemm_rs: 32..39)


@dataclass
class BatchedFit:
    beta: torch.Tensor          # [B, p]
    iterations: int
    objective: torch.Tensor     # [B] at the last check
    delta: torch.Tensor         # [B] largest coefficient change of the last checked iteration
    step: float
    history: list


class BatchedLassoLarge:
    """B lanes of the sample-sharded proximal-gradient solver advancing together - lane = (observation weights, lambda): the
    bootstrap replicates x lambda grid of a stability selection, or the lambda path of one fit, on a cohort that is
    streamed rather than cached.  With B in the hundreds both products of an iteration are tensor-core GEMMs:

        Eta^T = Vhl . X_r^T            ``ops.gemm.gemm_tn``  (tcgen05, bf16 hi / lo rows of v -> fp32-accurate eta)
        Rt    = split(W o (y - sigmoid(Eta)))                ``cml_batched_residual``
        G_own = sum_ranks Xt_r . Rt_r^T [owned genes]        ``XtrGemmReduceScatter`` (tcgen05 GEMM + reduce-scatter, one kernel)
        prox step on the owned genes, new Vhl rows stored to every rank   ``cml_batched_prox_allgather``

    The iteration is split into ``phase_a`` (local products, gradient launch) and ``phase_b`` (wait, proximal step, all-gather)
    so that several ranks can be played in one process (``symm.LocalGroup``: host execution model, single-GPU tests);
    ``run`` is the normal driver.  ``backend="unfused"`` is the torch / ``torch.distributed`` comparator and CPU path."""

    def __init__(self, x_local: torch.Tensor, y_local: torch.Tensor, lams, *, weights: Optional[torch.Tensor] = None,
                 n_total: Optional[int] = None, alpha: float = 1.0, step: Optional[float] = None, use_tma_reduce: bool = False,
                 rank: Optional[int] = None, world: Optional[int] = None, shared: Optional[dict] = None, prepared=None):
        """``prepared=(xb, xt)``: the two bf16 layouts of this rank's rows already on the device ([n_r, pad8(p)] sample-major and
        [p, pad8(n_r)] gene-major) - then ``x_local`` may be None and ``step`` must be given."""
        self.world = world if world is not None else (dist.get_world_size() if dist.is_initialized() else 1)
        self.rank = rank if rank is not None else (dist.get_rank() if dist.is_initialized() else 0)
        if prepared is not None:
            dev, n_r, p = prepared[0].device, prepared[0].shape[0], prepared[1].shape[0]
        else:
            dev = x_local.device
            n_r, p = x_local.shape
        self.n_r, self.p, self.dev, self.alpha = n_r, p, dev, float(alpha)
        lams = torch.as_tensor(np.asarray(lams, dtype=np.float32)).reshape(-1)
        self.B = int(lams.numel())
        self.Bp = (self.B + 31) // 32 * 32
        if n_total is None:
            t = torch.tensor([n_r], dtype=torch.float64, device=dev)
            all_reduce_(t)
            n_total = int(t.item())
        self.n_total = n_total
        if step is None and prepared is not None:
            raise ValueError("prepared operands need an explicit step")
        self.step = float(step) if step is not None else 1.0 / max(lipschitz_estimate(x_local, n_total), 1e-12)
        lam_pad = torch.full((self.Bp,), float("inf"))
        lam_pad[: self.B] = lams * self.alpha                     # padding lanes: infinite threshold, coefficients stay zero
        self.lam_l1 = lam_pad.to(dev)
        self.lams = lams.to(dev)
        self.shrink = 1.0                                         # per-lane ridge would need a vector; one alpha for all lanes
        if self.alpha < 1.0:
            if float(lams.max() - lams.min()) != 0.0:
                raise ValueError("elastic-net lanes must share one lambda (the ridge shrink factor is a scalar of the kernel)")
            self.shrink = 1.0 / (1.0 + self.step * float(lams[0]) * (1.0 - self.alpha))
        if prepared is not None:
            self.xb, self.xt = prepared
        else:
            xf = x_local.float()
            self.xb = F.to_bf16_padded(xf)                        # [n_r, ldp] sample-major: B operand of the eta product
            self.xt = F.to_bf16_padded(xf.t().contiguous())       # [p, ldn] gene-major: A operand of the gradient product
        self.ldp, self.ldn = self.xb.shape[1], self.xt.shape[1]
        self.y = y_local.float().contiguous()
        self.W = None
        if weights is not None:
            assert weights.shape == (self.B, n_r)
            self.W = torch.ones(self.Bp, n_r, dtype=torch.float32, device=dev)
            self.W[: self.B] = weights.float()
        gpr = F.XtrGemmReduceScatter.genes_per_rank_for(p, self.world)
        if shared is None:
            self.vhl = symm.alloc((2 * self.Bp, self.ldp), torch.bfloat16, dev)
            rs_bufs = None
        else:
            self.vhl, rs_bufs = shared["vhl"][self.rank], [b[self.rank] for b in shared["rs"]]
        self.rs = F.XtrGemmReduceScatter(p, self.Bp, dev, use_tma_reduce=use_tma_reduce, rank=self.rank, world=self.world, bufs=rs_bufs)
        go = self.rs.genes_owned
        self.ldo = max(go, 1)
        self.beta = torch.zeros(self.Bp, self.ldo, dtype=torch.float32, device=dev)
        self.v = torch.zeros(self.Bp, self.ldo, dtype=torch.float32, device=dev)
        self.lde = (n_r + 3) // 4 * 4
        self.eta2 = torch.empty(2 * self.Bp, self.lde, dtype=torch.float32, device=dev)
        self.rt = torch.zeros(2 * self.Bp, self.ldn, dtype=torch.bfloat16, device=dev)
        self.loss = torch.zeros(self.Bp, dtype=torch.float64, device=dev)
        self.dmax = torch.zeros(self.Bp, dtype=torch.float32, device=dev)
        self.done = torch.zeros(4, dtype=torch.int32, device=dev)
        self.tk = 1.0
        self.it = 0

    @staticmethod
    def shared_buffers(group: "symm.LocalGroup", p: int, n_lanes: int) -> dict:
        """The symmetric buffers of a ``LocalGroup`` run (one entry per rank in every list)."""
        Bp = (n_lanes + 31) // 32 * 32
        ldp = (p + 7) // 8 * 8
        gpr = F.XtrGemmReduceScatter.genes_per_rank_for(p, group.world)
        return {"vhl": group.alloc((2 * Bp, ldp), torch.bfloat16), "rs": [group.alloc((gpr, Bp), torch.float32) for _ in range(2)]}

    # ---- one iteration, in the two halves between which the other ranks have to have caught up
    def phase_a(self, check: bool) -> None:
        lb, P, st = _native.cuda(), _native.ptr, _native.stream_ptr
        self.it += 1
        if self.it > 1:      # every owner's new rows of v have landed in this rank's Vhl
            _native.check(lb.cml_wait_signals(ctypes.c_void_p(self.vhl.signal_ptrs[self.rank]), SIG_VHL, self.world, self.it - 1, st()),
                          "cml_wait_signals")
        from ..ops import gemm as G
        G.gemm_tn(self.vhl.local, [self.xb], self.n_r, out=self.eta2)
        if check:
            self.loss.zero_()
            self.dmax.zero_()
        _native.check(lb.cml_batched_residual(P(self.eta2), self.lde, self.Bp, self.n_r, P(self.y), P(self.W) if self.W is not None else None,
                                              self.n_r, P(self.rt), self.ldn, P(self.loss) if check else None, st()), "cml_batched_residual")
        self.rs.launch(self.xt, self.rt)

    def phase_b(self) -> None:
        lb, P, st = _native.cuda(), _native.ptr, _native.stream_ptr
        self.rs.wait()
        tk1 = (1.0 + np.sqrt(1.0 + 4.0 * self.tk * self.tk)) / 2.0
        mom = (self.tk - 1.0) / tk1
        self.tk = tk1
        grad = self.rs.bufs[self.rs.epoch & 1].local
        _native.check(lb.cml_batched_prox_allgather(P(grad), self.Bp, self.rs.genes_owned, self.rs.gene0, self.Bp, P(self.lam_l1), self.step,
                                                    1.0 / self.n_total, mom, self.shrink, P(self.beta), P(self.v), self.ldo,
                                                    F._ptr_array(self.vhl.peer_ptrs), self.ldp, F._ptr_array(self.vhl.signal_ptrs), self.rank,
                                                    self.world, P(self.done), self.it, SIG_VHL, P(self.dmax), st()),
                      "cml_batched_prox_allgather")

    def wait_all_gather(self) -> None:
        _native.check(_native.cuda().cml_wait_signals(ctypes.c_void_p(self.vhl.signal_ptrs[self.rank]), SIG_VHL, self.world, self.it,
                                                      _native.stream_ptr()), "cml_wait_signals")

    def penalty_owned(self) -> torch.Tensor:
        """[B]: lambda (alpha |b|_1 + (1 - alpha)/2 |b|^2) over the genes this rank owns."""
        b = self.beta[: self.B, : self.rs.genes_owned].double()
        return self.lams.double() * (self.alpha * b.abs().sum(1) + 0.5 * (1.0 - self.alpha) * (b * b).sum(1))

    def coefficients(self) -> torch.Tensor:
        """[B, p] on every rank (one all-gather of the owned slices)."""
        own = torch.zeros(self.B, self.rs.genes_per_rank, dtype=torch.float32, device=self.dev)
        own[:, : self.rs.genes_owned] = self.beta[: self.B, : self.rs.genes_owned]
        if self.world > 1 and dist.is_initialized():
            parts = [torch.empty_like(own) for _ in range(self.world)]
            dist.all_gather(parts, own)
            own = torch.cat(parts, 1)
        return own[:, : self.p].contiguous()

    def run(self, max_iter: int = 200, tol: float = 1e-5, check_every: int = 10) -> BatchedFit:
        history = []
        obj = torch.full((self.B,), float("nan"), dtype=torch.float64)
        d = torch.full((self.B,), float("inf"))
        for it in range(1, max_iter + 1):
            check = (it % check_every == 0) or it == max_iter
            self.phase_a(check)
            self.phase_b()
            if check:
                dm = self.dmax[: self.B].double().clone()
                s = self.loss[: self.B].clone() / self.n_total + self.penalty_owned()
                all_reduce_(dm, dist.ReduceOp.MAX)
                all_reduce_(s)
                obj, d = s.cpu(), dm.float().cpu()
                history.append((it, float(obj.mean()), float(d.max())))
                if float(d.max()) < tol:
                    break
        self.wait_all_gather()
        _native.synchronize()
        self.vhl.barrier()
        return BatchedFit(self.coefficients(), self.it, obj, d, self.step, history)


def fit_batched(x_local: torch.Tensor, y_local: torch.Tensor, lams, *, weights: Optional[torch.Tensor] = None, n_total: Optional[int] = None,
                max_iter: int = 200, tol: float = 1e-5, backend: str = "auto", alpha: float = 1.0, step: Optional[float] = None,
                check_every: int = 10, use_tma_reduce: bool = False) -> BatchedFit:
    """All lanes ``(weights[b], lams[b])`` of the sample-sharded lasso at once; every rank returns ``beta`` [B, p].
    ``backend``: "fused" (tensor-core products, in-kernel reduce-scatter and all-gather; CUDA or the host execution model),
    "unfused" (torch matmul + ``torch.distributed`` all-reduce: the comparator and the plain CPU path)."""
    if backend == "auto":
        backend = "fused" if _native.on_device(x_local) else "unfused"
    if backend == "fused":
        return BatchedLassoLarge(x_local, y_local, lams, weights=weights, n_total=n_total, alpha=alpha, step=step,
                                 use_tma_reduce=use_tma_reduce).run(max_iter, tol, check_every)
    dev = x_local.device
    lam = torch.as_tensor(np.asarray(lams, dtype=np.float32)).reshape(-1).to(dev)
    B, (n_r, p) = int(lam.numel()), x_local.shape
    if n_total is None:
        t = torch.tensor([n_r], dtype=torch.float64, device=dev)
        all_reduce_(t)
        n_total = int(t.item())
    if step is None:
        step = 1.0 / max(lipschitz_estimate(x_local, n_total), 1e-12)
    xf, yf = x_local.float(), y_local.float()
    W = weights.float() if weights is not None else torch.ones(B, n_r, device=dev)
    beta = torch.zeros(B, p, device=dev)
    v = beta.clone()
    tk, history = 1.0, []
    obj = torch.full((B,), float("nan"), dtype=torch.float64)
    d = torch.full((B,), float("inf"))
    it = 0
    shrink = 1.0 / (1.0 + step * lam * (1.0 - alpha))
    for it in range(1, max_iter + 1):
        eta = v @ xf.t()                                          # [B, n_r]
        r = W * (yf[None, :] - torch.sigmoid(eta))
        g = r @ xf                                                # [B, p]
        all_reduce_(g)
        z = v + step * g / n_total
        thr = (step * lam * alpha)[:, None]
        bnew = torch.sign(z) * (z.abs() - thr).clamp_min(0) * shrink[:, None]
        tk1 = (1.0 + np.sqrt(1.0 + 4.0 * tk * tk)) / 2.0
        mom = (tk - 1.0) / tk1
        tk = tk1
        dd = (bnew - beta).abs().amax(1)
        v = bnew + mom * (bnew - beta)
        beta = bnew
        if it % check_every == 0 or it == max_iter:
            sp = torch.where(eta > 0, eta + torch.log1p(torch.exp(-eta)), torch.log1p(torch.exp(eta)))
            s = (W * (sp - yf[None, :] * eta)).double().sum(1)
            all_reduce_(s)
            bd = beta.double()
            obj = (s / n_total + lam.double() * (alpha * bd.abs().sum(1) + 0.5 * (1.0 - alpha) * (bd * bd).sum(1))).cpu()
            d = dd.cpu()
            history.append((it, float(obj.mean()), float(d.max())))
            if float(d.max()) < tol:
                break
    return BatchedFit(beta, it, obj, d, step, history)


def stability_selection_large(x_local: torch.Tensor, y_local: torch.Tensor, lams, n_boot: int = 64, *, row0: Optional[int] = None,
                              n_total: Optional[int] = None, seed: int = 0, max_iter: int = 150, tol: float = 1e-4, backend: str = "auto",
                              alpha: float = 1.0, lanes_per_batch: int = 256, threshold: float = 1e-6) -> dict:
    """Stability selection on a sample-sharded cohort: ``n_boot`` bootstrap replicates x the lambda grid, all (replicate, lambda)
    pairs fitted as lanes of the batched solver, ``lanes_per_batch`` at a time.  The bootstrap is the Poisson(1) weighting of the
    samples, drawn per replicate from a generator keyed by (seed, replicate) over the GLOBAL sample index, so every sharding of
    the cohort sees the same replicates.  Returns the selection frequency of every gene at every lambda ([n_lambda, p]) and
    its maximum over the grid (the reference's bootstraps x alpha grid, models.lasso.stability_selection, at cohort sizes where
    the design matrix is streamed)."""
    lams = np.asarray(lams, dtype=np.float64).reshape(-1)
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    n_r, p = x_local.shape
    dev = x_local.device
    if n_total is None or row0 is None:
        counts = torch.zeros(world, dtype=torch.float64, device=dev)
        counts[rank] = n_r
        all_reduce_(counts)
        n_total = int(counts.sum().item())
        row0 = int(counts[:rank].sum().item())
    W = torch.empty(n_boot, n_r, dtype=torch.float32)
    for b in range(n_boot):
        g = torch.Generator().manual_seed(seed * 1000003 + b)
        W[b] = torch.poisson(torch.ones(n_total), generator=g)[row0: row0 + n_r]
    W = W.to(dev)
    step = 1.0 / max(lipschitz_estimate(x_local, n_total), 1e-12) / 2.0          # bootstrap weights up to ~2x the curvature bound
    pairs = [(b, l) for b in range(n_boot) for l in range(len(lams))]
    freq = torch.zeros(len(lams), p, dtype=torch.float64, device=dev)
    iters = []
    for s in range(0, len(pairs), lanes_per_batch):
        chunk = pairs[s: s + lanes_per_batch]
        wts = W[[b for b, _ in chunk]]
        ls = lams[[l for _, l in chunk]]
        fit = fit_batched(x_local, y_local, ls, weights=wts, n_total=n_total, max_iter=max_iter, tol=tol, backend=backend, alpha=alpha, step=step)
        sel = (fit.beta.abs() > threshold).double()
        idx = torch.as_tensor([l for _, l in chunk], device=dev)
        freq.index_add_(0, idx, sel)
        iters.append(fit.iterations)
    freq /= n_boot
    return {"frequency": freq, "max_frequency": freq.max(0).values, "lambdas": lams, "iterations": iters, "step": step, "n_total": n_total}
