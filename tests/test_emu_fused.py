"""The fused compute + collective kernels with SEVERAL ranks, on the CPU.

Under the host execution model (csrc/emu/cuda_emu.h) a "rank" is just a set of buffers: the kernels take tables of peer
pointers, so the ranks of a node can be played one after the other in one process - every rank's kernel adds into / stores to
the other ranks' buffers exactly as it would over NVLink, then the wait kernels check the signal words.  (SURVEY.md section 4:
"single-GPU multi-rank emulation"; here it needs no GPU at all.)  The comparators are the plain torch computations that the
NCCL-based unfused paths perform."""
import ctypes

import numpy as np
import pytest
import torch

from consensusml_b200 import _native
from consensusml_b200.parallel import fused as F      # registers the signatures

P = _native.ptr


def _ptrs(ts):
    return (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])


@pytest.mark.parametrize("world", [1, 2, 3])
def test_histogram_reduce_scatter_and_allreduce_landings(world):
    g = torch.Generator().manual_seed(world)
    n_loc, p, nbins, n_nodes = 300, 21, 16, 4
    fpr = ((p + world - 1) // world + 7) // 8 * 8
    bins = [torch.randint(0, nbins, (p, 304), generator=g, dtype=torch.uint8) for _ in range(world)]
    node = [torch.randint(-1, n_nodes, (n_loc,), generator=g).to(torch.int16) for _ in range(world)]
    gh = [torch.stack([torch.rand(n_loc, generator=g) * 2 - 1, torch.rand(n_loc, generator=g) * 0.25], 1).contiguous() for _ in range(world)]
    # only nodes 1 and 2 are built, into slots 0 and 1 (sibling subtraction would derive the other two)
    slot_of = torch.tensor([-1, 0, 1, -1], dtype=torch.int16)
    n_slots = 2
    expect = torch.zeros(n_slots, p, nbins, 2, dtype=torch.float64)
    for r in range(world):
        for i in range(n_loc):
            s = int(slot_of[node[r][i]]) if node[r][i] >= 0 else -1
            if s >= 0:
                for f in range(p):
                    expect[s, f, int(bins[r][f, i])] += gh[r][i].double()
    with _native.emulation(seed=7, sms=2):
        lib = _native.cuda()
        # ---- reduce-scatter landing: rank r receives features [r * fpr, (r + 1) * fpr)
        landing = [torch.zeros(n_slots * fpr * nbins * 2) for _ in range(world)]
        sig = [torch.zeros(64, dtype=torch.int32) for _ in range(world)]
        for r in range(world):
            local = torch.zeros(n_slots * p * nbins * 2)
            cnt = torch.zeros(p + 16, dtype=torch.int32)
            rc = lib.cml_hist_build(P(bins[r]), 304, n_loc, p, nbins, P(node[r]), P(gh[r]), n_slots, P(slot_of), P(local), P(cnt), None,
                                    _ptrs(landing), _ptrs(sig), r, world, fpr, 1, F.SIG_HIST, 1.0, 0.25, None)
            assert rc == 0, lib.cml_last_error()
            assert float(local.abs().max()) == 0.0 and int(cnt.abs().max()) == 0           # scratch is self-cleaning
        for r in range(world):
            assert lib.cml_wait_signals(P(sig[r]), F.SIG_HIST, world, 1, None) == 0         # every rank has signalled this owner
            got = landing[r].view(n_slots, fpr, nbins, 2)
            lo, hi = r * fpr, min(p, (r + 1) * fpr)
            if hi > lo:
                assert torch.allclose(got[:, : hi - lo].double(), expect[:, lo:hi], atol=2e-5), r
            assert float(got[:, max(hi - lo, 0):].abs().max() if hi - lo < fpr else 0.0) == 0.0
        # ---- all-reduce landing (peer red.add to every rank): every rank ends up with the whole histogram
        full = [torch.zeros(n_slots * p * nbins * 2) for _ in range(world)]
        for r in range(world):
            local = torch.zeros(n_slots * p * nbins * 2)
            cnt = torch.zeros(p + 16, dtype=torch.int32)
            rc = lib.cml_hist_build(P(bins[r]), 304, n_loc, p, nbins, P(node[r]), P(gh[r]), n_slots, P(slot_of), P(local), P(cnt), None,
                                    _ptrs(full), _ptrs(sig), r, world, 0, 2, F.SIG_HIST, 1.0, 0.25, None)
            assert rc == 0
        for r in range(world):
            assert lib.cml_wait_signals(P(sig[r]), F.SIG_HIST, world, 2, None) == 0
            assert torch.allclose(full[r].view(n_slots, p, nbins, 2).double(), expect, atol=2e-5)
            assert torch.equal(full[r], full[0])


@pytest.mark.parametrize("world,nb", [(1, 1), (2, 4), (3, 8)])
def test_xtr_reduce_scatter_then_proximal_step_and_allgather(world, nb):
    g = torch.Generator().manual_seed(10 + world)
    n_loc, p = 260, 1100                                   # two gene tiles of 1024
    gpr = (p + world - 1) // world
    ld = (p + 7) // 8 * 8
    X = [torch.randn(n_loc, p, generator=g) for _ in range(world)]
    Xb = []
    for r in range(world):
        t = torch.zeros(n_loc, ld, dtype=torch.bfloat16)
        t[:, :p] = X[r].to(torch.bfloat16)
        Xb.append(t)
    R = [torch.randn(n_loc, nb, generator=g).contiguous() for _ in range(world)]
    expect = sum(Xb[r][:, :p].float().t().double() @ R[r].double() for r in range(world))        # [p, nb]
    with _native.emulation(seed=3, sms=2):
        lib = _native.cuda()
        grad = [torch.zeros(gpr * nb) for _ in range(world)]
        sig = [torch.zeros(64, dtype=torch.int32) for _ in range(world)]
        for r in range(world):
            scratch = torch.zeros(lib.cml_xtr_rs_scratch_bytes(p, nb), dtype=torch.uint8)
            rc = lib.cml_xtr_reduce_scatter(P(Xb[r]), ld, n_loc, p, P(R[r]), nb, P(scratch), _ptrs(grad), _ptrs(sig), r, world, gpr, 0, 1, None)
            assert rc == 0, lib.cml_last_error()
            assert int(scratch.abs().max()) == 0
        for r in range(world):
            assert lib.cml_wait_signals(P(sig[r]), F.SIG_GRAD, world, 1, None) == 0
            lo, hi = r * gpr, min(p, (r + 1) * gpr)
            got = grad[r][: (hi - lo) * nb].view(hi - lo, nb).double()
            assert torch.allclose(got, expect[lo:hi], rtol=1e-4, atol=1e-3), r
        if nb != 1:
            return
        # owner-side proximal step fused with the all-gather of the new coefficients (peer stores into every rank's copy)
        beta = [torch.zeros(p) for _ in range(world)]
        v = [torch.zeros(p) for _ in range(world)]
        dmax = [torch.zeros(1) for _ in range(world)]
        step, lam, inv_n = 0.5, 0.05, 1.0 / (n_loc * world)
        for r in range(world):
            lo, hi = r * gpr, min(p, (r + 1) * gpr)
            done = torch.zeros(4, dtype=torch.int32)
            rc = lib.cml_prox_allgather(P(grad[r]), hi - lo, lo, None, step, lam, inv_n, 0.0, _ptrs(beta), _ptrs(v), _ptrs(sig), r, world,
                                        P(beta[r]), P(done), 1, F.SIG_BETA, P(dmax[r]), None)
            assert rc == 0
        z = step * expect[:, 0] * inv_n
        ref = torch.sign(z) * (z.abs() - step * lam).clamp_min(0)
        for r in range(world):
            assert lib.cml_wait_signals(P(sig[r]), F.SIG_BETA, world, 1, None) == 0
            assert torch.allclose(beta[r].double(), ref, atol=1e-5) and torch.equal(beta[r], beta[0]) and torch.equal(v[r], beta[r])
            assert float(grad[r].abs().max()) == 0.0                                        # cleared for the next epoch


@pytest.mark.parametrize("world", [1, 3])
def test_shard_transpose_kernel(world):
    g = torch.Generator().manual_seed(5)
    rows_loc, cols = 70, 45
    x = [torch.randn(rows_loc, cols, generator=g) for _ in range(world)]
    full = torch.cat(x)                                     # [rows, cols]
    cs = [cols * r // world for r in range(world + 1)]
    ldy = rows_loc * world
    with _native.emulation(seed=2):
        lib = _native.cuda()
        lib.cml_shard_transpose.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.c_uint, ctypes.c_int, ctypes.c_void_p]
        out = [torch.zeros(max(1, cs[r + 1] - cs[r]), ldy) for r in range(world)]
        sig = [torch.zeros(64, dtype=torch.int32) for _ in range(world)]
        cs_t = torch.tensor(cs, dtype=torch.int32)
        for r in range(world):
            done = torch.zeros(4, dtype=torch.int32)
            rc = lib.cml_shard_transpose(P(x[r]), cols, rows_loc, cols, P(cs_t), _ptrs(out), ldy, r * rows_loc, _ptrs(sig), r, world, P(done), 1, 24, None)
            assert rc == 0, lib.cml_last_error()
        for r in range(world):
            assert lib.cml_wait_signals(P(sig[r]), 24, world, 1, None) == 0
            assert torch.equal(out[r][: cs[r + 1] - cs[r]], full[:, cs[r]:cs[r + 1]].t())


def test_xb_residual_and_split_scoring_kernels():
    g = torch.Generator().manual_seed(8)
    n, p = 90, 40
    ld = 40
    x = torch.randn(n, p, generator=g).to(torch.bfloat16)
    v = torch.randn(p, generator=g) * 0.3
    y = (torch.rand(n, generator=g) < 0.5).float()
    w = torch.rand(n, generator=g)
    R, eta, loss = torch.zeros(n), torch.zeros(n), torch.zeros(1, dtype=torch.float64)
    hist = torch.rand(3, 7, 16, 2, generator=g)
    hist[..., 0] -= 0.5
    hist[1, 2] = 0                                          # an empty feature
    with _native.emulation(seed=1, sms=2):
        lib = _native.cuda()
        rc = lib.cml_xb_residual(P(x), ld, n, p, P(v), ctypes.c_float(0.2), P(y), P(w), P(R), P(eta), P(loss), None)
        assert rc == 0, lib.cml_last_error()
        gain, feat, thr = F.best_splits(hist, 1.0, 0.0, 0.1)
    e = x.float() @ v + 0.2
    q = torch.sigmoid(e)
    assert torch.allclose(eta, e, atol=1e-5) and torch.allclose(R, w * (y - q), atol=1e-5)
    ll = (w.double() * (torch.nn.functional.softplus(e.double()) - y.double() * e.double())).sum()
    assert abs(float(loss) - float(ll)) < 1e-4
    g2, f2, t2 = F.best_splits_torch(hist, 1.0, 0.0, 0.1)
    assert torch.equal(feat, f2) and torch.equal(thr, t2) and torch.allclose(gain, g2, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("world", [1, 3])
def test_deterministic_reduce_scatter_is_bit_reproducible_and_exact(world):
    """The fixed-order form of X^T R reduce-scatter: chunk slices added in chunk order, per-source slots added in rank order."""
    g = torch.Generator().manual_seed(30 + world)
    n_loc, p, nb = 700, 1100, 4
    gpr = (p + world - 1) // world
    ld = (p + 7) // 8 * 8
    Xb = []
    for r in range(world):
        t = torch.zeros(n_loc, ld, dtype=torch.bfloat16)
        t[:, :p] = torch.randn(n_loc, p, generator=g).to(torch.bfloat16)
        Xb.append(t)
    R = [torch.randn(n_loc, nb, generator=g).contiguous() for _ in range(world)]
    expect = sum(Xb[r][:, :p].float().t().double() @ R[r].double() for r in range(world))
    runs = []
    for seed in (1, 9):
        with _native.emulation(seed=seed, sms=2):
            lib = _native.cuda()
            slots = [torch.full((world * gpr * nb,), float("nan")) for _ in range(world)]        # stale contents must not matter
            sig = [torch.zeros(64, dtype=torch.int32) for _ in range(world)]
            for r in range(world):
                scratch = torch.zeros(lib.cml_xtr_rs_det_scratch_bytes(n_loc, p, nb), dtype=torch.uint8)
                rc = lib.cml_xtr_reduce_scatter_det(P(Xb[r]), ld, n_loc, p, P(R[r]), nb, P(scratch), _ptrs(slots), _ptrs(sig), r, world, gpr, 1, None)
                assert rc == 0, lib.cml_last_error()
            grads = []
            for r in range(world):
                assert lib.cml_wait_signals(P(sig[r]), F.SIG_GRAD, world, 1, None) == 0
                gr = torch.empty(gpr * nb)
                lo, hi = r * gpr, min(p, (r + 1) * gpr)
                # genes beyond p in the last owner's slice were never written: reduce only what exists
                assert lib.cml_xtr_det_reduce(P(slots[r]), P(gr), gpr, nb, world, None) == 0
                grads.append(gr[: (hi - lo) * nb].view(hi - lo, nb).clone())
                assert torch.allclose(grads[-1].double(), expect[lo:hi], rtol=1e-4, atol=1e-3)
        runs.append(grads)
    for a, b in zip(*runs):
        assert torch.equal(a, b)                     # same bits whatever the thread order
