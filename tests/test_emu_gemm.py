"""The tcgen05 / TMA kernels on the host model of the asynchronous units (csrc/emu/async_emu.cpp).

gemm_tcgen05.cu is the calibration: it ran on hardware in round 1, so a model that disagreed with the hardware about
descriptors, the 128B swizzle, TMEM layout or mbarrier phases would fail here.  gemm_rs.cu has never run on hardware; the same
model executes it below (register-reduce and TMA-reduce landing, hi / lo split, split-K, several owners and ranks).
Asynchronous operations complete as late as possible and in a seed-dependent order, so a missing wait shows up as stale or
poisoned data: the last test removes single waits from the kernel and checks that the result breaks."""
import ctypes
import os
import subprocess
import sys

import pytest
import torch

from consensusml_b200 import _native
from consensusml_b200.ops import gemm as G
from consensusml_b200.parallel import fused as F


def _bf16_pad(x):
    rows, k = x.shape
    ld = (k + 7) // 8 * 8
    y = torch.zeros(rows, ld, dtype=torch.bfloat16)
    y[:, :k] = x.bfloat16()
    return y


@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("M,N,K,seed", [(100, 90, 128, 0), (300, 520, 200, 3), (257, 769, 72, 11), (130, 1000, 328, 5)])
def test_gemm_epilogues(variant, M, N, K, seed):
    torch.manual_seed(seed)
    x, z = torch.randn(M, K), torch.randn(N, K)
    with _native.emulation(seed=seed):
        G.set_variant(variant)
        try:
            xa, na = G.to_bf16_rows(x)
            xb, nb = G.to_bf16_rows(z)
            c = G.gemm_tn(xa, [xb], N, epilogue=G.EPI_SCALE, alpha=0.5)
            d = G.gemm_tn(xa, [xb], N, epilogue=G.EPI_SQDIST, norm_a=na, norm_b=[nb])
            r = G.gemm_tn(xa, [xb], N, epilogue=G.EPI_RBF, alpha=1.0 / K, norm_a=na, norm_b=[nb])
        finally:
            G.set_variant(0)
    prod = xa.float() @ xb.float().t()
    d2 = (na[:, None] + nb[None, :] - 2 * prod).clamp_min(0)
    assert (c - 0.5 * prod).abs().max() < 1e-4
    assert (d - d2).abs().max() < 2e-4
    assert (r - torch.exp(-d2 / K)).abs().max() < 1e-5
    assert (na - xa.float().square().sum(1)).abs().max() < 1e-3


def test_gemm_unaligned_output_takes_register_stores():
    """A C whose pitch TMA cannot address (ldc % 4 != 0 is impossible through gemm_tn, so: an out view offset by one float)."""
    torch.manual_seed(2)
    x, z = torch.randn(140, 96), torch.randn(75, 96)
    with _native.emulation(seed=4):
        xa, _ = G.to_bf16_rows(x)
        xb, _ = G.to_bf16_rows(z)
        buf = torch.zeros(140 * 76 + 1)
        out = buf[1:].view(140, 76)                       # 4-byte aligned only
        c = G.gemm_tn(xa, [xb], 75, out=out)
    assert (c - xa.float() @ xb.float().t()).abs().max() < 1e-4


def test_gemm_many_sources_direct():
    """B spread over three row blocks (the direct all-gather form: one tensor map per source, ragged last block)."""
    torch.manual_seed(3)
    x = torch.randn(150, 136)
    blocks = [torch.randn(200, 136), torch.randn(200, 136), torch.randn(117, 136)]
    with _native.emulation(seed=6):
        xa, na = G.to_bf16_rows(x)
        bs = [G.to_bf16_rows(b) for b in blocks]
        c = G.gemm_tn(xa, [b[0] for b in bs], 517, epilogue=G.EPI_SQDIST, norm_a=na, norm_b=[b[1] for b in bs])
    zb = torch.cat([b[0].float() for b in bs])
    nb = torch.cat([b[1] for b in bs])
    ref = (na[:, None] + nb[None, :] - 2 * xa.float() @ zb.t()).clamp_min(0)
    assert (c - ref).abs().max() < 2e-4


def test_gemm_staged_unaligned_source_blocks():
    """8 sources of 750 rows: output column starts that are not 16-byte aligned (3000 B).  This is the shape whose TMA-store
    epilogue faulted on the 8-GPU run of round 1; the chunks that cannot be TMA-stored must take the register path (the
    model stops on a TMA store with an unaligned innermost coordinate, as the hardware did)."""
    torch.manual_seed(8)
    M, K, rows, n_src = 140, 72, 750, 8
    N = n_src * rows
    with _native.emulation(seed=5, sms=8):
        xa, na = G.to_bf16_rows(torch.randn(M, K))
        srcs = [G.to_bf16_rows(torch.randn(rows, K)) for _ in range(n_src)]
        ldb = srcs[0][0].shape[1]
        stage = torch.zeros(n_src * rows, ldb, dtype=torch.bfloat16)
        stage_norm = torch.zeros(n_src * rows)
        flags = torch.zeros(_native.cuda().cml_gemm_staged_flags(n_src, rows), dtype=torch.int32)
        c = G.gemm_tn_staged(xa, [s[0].data_ptr() for s in srcs], N, rows, ldb, stage, stage_norm, flags, 1, 3, epilogue=G.EPI_SQDIST,
                             norm_a=na, peer_norm_ptrs=[s[1].data_ptr() for s in srcs], n_copiers=2)
    zb = torch.cat([s[0].float() for s in srcs])
    nb = torch.cat([s[1] for s in srcs])
    assert (c - (na[:, None] + nb[None, :] - 2 * xa.float() @ zb.t()).clamp_min(0)).abs().max() < 2e-4


@pytest.mark.parametrize("variant,n_src,rank,grouped", [(1, 2, 0, 0), (2, 3, 1, 0), (2, 2, 1, 1), (1, 3, 2, 1)])
def test_gemm_staged_all_gather(variant, n_src, rank, grouped):
    """Copier CTAs pull the source blocks into the staging buffer and flag tiles; GEMM CTAs of the same launch consume them."""
    torch.manual_seed(10 + n_src)
    M, K, rows = 260, 200, 300
    N = n_src * rows - 37                                   # ragged last source
    x = torch.randn(M, K)
    with _native.emulation(seed=variant + rank, sms=8):
        G.set_variant(variant)
        G.set_staged_grouping(bool(grouped))
        try:
            xa, na = G.to_bf16_rows(x)
            srcs = [G.to_bf16_rows(torch.randn(rows, K)) for _ in range(n_src)]
            ldb = srcs[0][0].shape[1]
            stage = torch.full((n_src * rows, ldb), float("nan"), dtype=torch.bfloat16)
            stage_norm = torch.full((n_src * rows,), float("nan"))
            n_flags = _native.cuda().cml_gemm_staged_flags(n_src, rows)
            flags = torch.zeros(n_flags, dtype=torch.int32)
            for epoch in (1, 2):                              # flags are compared for equality: the second call must not reuse the first's
                if epoch == 2:
                    stage.fill_(float("nan"))
                c = G.gemm_tn_staged(xa, [s[0].data_ptr() for s in srcs], N, rows, ldb, stage, stage_norm, flags, epoch, rank,
                                     epilogue=G.EPI_SQDIST, norm_a=na, peer_norm_ptrs=[s[1].data_ptr() for s in srcs], n_copiers=2).clone()
                zb = torch.cat([s[0].float() for s in srcs])[:N]
                nb = torch.cat([s[1] for s in srcs])[:N]
                ref = (na[:, None] + nb[None, :] - 2 * xa.float() @ zb.t()).clamp_min(0)
                assert (c - ref).abs().max() < 2e-4
                assert bool((flags == epoch).all())
        finally:
            G.set_variant(0)
            G.set_staged_grouping(False)


def _run_gemm_rs(p, B, K, world, use_tma, split, seed, sms=None):
    g = torch.Generator().manual_seed(seed + 1)
    xts = [_bf16_pad(torch.randn(p, K, generator=g)) for _ in range(world)]
    rfs = [torch.randn(B, K, generator=g) for _ in range(world)]
    ld = xts[0].shape[1]
    rts = []
    for rf in rfs:
        pad = torch.zeros(B, ld)
        pad[:, :K] = rf
        rts.append(F.split_bf16(pad) if split else pad.bfloat16())
    gpr = ((p + world - 1) // world + 31) // 32 * 32
    outs = [torch.zeros(gpr, B) for _ in range(world)]
    sigs = [torch.zeros(256, dtype=torch.int32) for _ in range(world)]
    arr = lambda ts: (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    with _native.emulation(seed=seed, sms=sms):
        for r in range(world):
            done = torch.zeros(4, dtype=torch.int32)
            rc = _native.cuda().cml_xtr_gemm_reduce_scatter(p, rts[r].shape[0], B, ld, _native.ptr(xts[r]), ld, _native.ptr(rts[r]), ld, arr(outs), B,
                                                            gpr, arr(sigs), r, world, _native.ptr(done), 7, F.SIG_GEMM_RS, 0.5, int(use_tma),
                                                            _native.stream_ptr())
            _native.check(rc, "cml_xtr_gemm_reduce_scatter")
            assert int(done[0]) == 0                          # the last CTA re-arms the counter
    ref = sum(0.5 * (xts[r].float() @ (rts[r].float()[:B] + (rts[r].float()[B:] if split else 0)).t()) for r in range(world))
    exact = sum(0.5 * (xts[r].float()[:, :K] @ rfs[r].t()) for r in range(world))
    return outs, sigs, ref, exact, gpr


@pytest.mark.parametrize("p,B,K,world,use_tma,split,seed,sms", [
    (512, 256, 300, 1, 0, 0, 0, None), (512, 256, 300, 1, 1, 0, 2, None), (1000, 128, 520, 3, 0, 0, 5, None),
    (1000, 128, 2100, 2, 1, 1, 9, None), (300, 64, 4000, 2, 0, 1, 0, 8), (777, 96, 1300, 4, 1, 0, 13, 6)])
def test_gemm_reduce_scatter(p, B, K, world, use_tma, split, seed, sms):
    outs, sigs, ref, exact, gpr = _run_gemm_rs(p, B, K, world, use_tma, split, seed, sms)
    scale = float(ref.abs().max())
    for d in range(world):
        lo, hi = min(p, d * gpr), min(p, (d + 1) * gpr)
        assert (outs[d][: hi - lo] - ref[lo:hi]).abs().max() < 2e-5 * scale          # same bf16 operands, fp32 accumulation
        assert float(outs[d][hi - lo:].abs().max()) == 0.0 if hi - lo < gpr else True   # rows past the owner's genes stay untouched
        assert [int(sigs[d][F.SIG_GEMM_RS + r]) for r in range(world)] == [7] * world     # every rank signalled every owner
        if split:                                                                         # hi + lo carries the fp32 residuals
            assert (outs[d][: hi - lo] - exact[lo:hi]).abs().max() < 3e-4 * scale


def test_gemm_eager_completion_mode(monkeypatch):
    monkeypatch.setenv("CML_EMU_ASYNC", "eager")
    outs, _, ref, _, gpr = _run_gemm_rs(512, 128, 300, 2, 1, 0, 1)
    assert (torch.cat(outs)[:512] - ref).abs().max() < 2e-5 * float(ref.abs().max())


def test_high_level_ops_use_the_kernel_under_emulation():
    torch.manual_seed(5)
    x, z = torch.randn(93, 200), torch.randn(44, 200)
    with _native.emulation(seed=1):
        _native.reset_calls()
        k = G.rbf_kernel(x, z)
        assert _native.calls.get("cml_gemm_tn_bf16", 0) == 1
    ref = torch.exp(-torch.cdist(x.double(), z.double()) ** 2 / 200).float()
    assert (k - ref).abs().max() < 5e-3                    # bf16 operands


_FAULT = r"""
import sys, torch
root = %r
sys.path[:0] = [root, root + "/tests"]
from test_emu_gemm import _run_gemm_rs
outs, _, ref, _, gpr = _run_gemm_rs(512, 256, 300, 1, %d, 0, 0)
bad = not bool(((outs[0][:512] - ref).abs().max() < 1e-3).item())
print("RESULT", "wrong" if bad else "right")
"""


@pytest.mark.parametrize("n,use_tma", [(1, 0), (300, 0), (560, 1), (800, 1)])
def test_a_removed_wait_is_detected(n, use_tma):
    """Remove the n-th mbarrier wait (these four are epilogue / MMA-issuer / producer waits that would have blocked): the
    late-completion model must turn that into a wrong result or a rule violation, never into a pass."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CML_EMU_FAULT=f"skip_wait:{n}", CML_EMU_FAULT_VERBOSE="1")
    res = subprocess.run([sys.executable, "-c", _FAULT % (root, use_tma)], env=env, capture_output=True, text=True, timeout=600, cwd=root)
    assert "skipped wait" in res.stderr
    line = [l for l in res.stderr.splitlines() if "skipped wait" in l][0]
    parity, phase = int(line.split("parity ")[1].split()[0]), int(line.split("phase now ")[1].split(")")[0])
    assert parity == phase, "test premise: the removed wait would have blocked"
    assert res.returncode != 0 or "RESULT wrong" in res.stdout


def _batched_problem(n, p, B, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, p, generator=g).bfloat16().float()          # bf16-exact design: both backends see the same numbers
    bt = torch.zeros(p)
    bt[:8] = torch.randn(8, generator=g) * 1.5
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ bt)).float()
    from consensusml_b200.models import lasso_large as LL
    lmax = LL.lambda_max(x, y, n)
    import numpy as np
    lams = np.geomspace(lmax * 0.9, lmax * 0.1, B)
    W = torch.poisson(torch.ones(B, n), generator=g)
    return x, y, lams, W


@pytest.mark.parametrize("use_tma", [False, True])
def test_batched_sharded_lasso_single_rank(use_tma):
    """The replicate-batched iteration end to end on the kernels: tcgen05 eta product, residual split, fused GEMM ->
    reduce-scatter gradient, proximal step + all-gather; against the torch iteration of the same lanes."""
    from consensusml_b200.models import lasso_large as LL
    x, y, lams, W = _batched_problem(150, 260, 20, 0)
    ref = LL.fit_batched(x, y, lams, weights=W, backend="unfused", max_iter=40, tol=0.0)
    with _native.emulation(seed=3):
        got = LL.BatchedLassoLarge(x, y, lams, weights=W, step=ref.step, use_tma_reduce=use_tma).run(40, 0.0, 10)
    assert got.iterations == ref.iterations == 40
    assert (got.beta - ref.beta).abs().max() < 2e-4
    assert (got.objective - ref.objective).abs().max() < 1e-5
    assert torch.equal((got.beta.abs() > 1e-4).sum(1), (ref.beta.abs() > 1e-4).sum(1))
    assert len(got.history) == 4 and got.history[-1][1] < got.history[0][1]


@pytest.mark.parametrize("world", [2, 3])
def test_batched_sharded_lasso_several_ranks_in_one_process(world):
    """Ranks played one after the other on the host model (symm.LocalGroup): every rank's gradient kernel reduces into the
    owners' buffers, every owner's proximal kernel stores its rows of v into all ranks' copies."""
    from consensusml_b200.models import lasso_large as LL
    from consensusml_b200.parallel import symm
    n, p, B = 210, 300, 12
    x, y, lams, W = _batched_problem(n, p, B, world)
    ref = LL.fit_batched(x, y, lams, weights=W, backend="unfused", max_iter=25, tol=0.0)
    bounds = [n * r // world for r in range(world + 1)]
    with _native.emulation(seed=world, sms=6):
        group = symm.LocalGroup(world)
        shared = LL.BatchedLassoLarge.shared_buffers(group, p, B)
        ranks = [LL.BatchedLassoLarge(x[bounds[r]:bounds[r + 1]], y[bounds[r]:bounds[r + 1]], lams, weights=W[:, bounds[r]:bounds[r + 1]],
                                      n_total=n, step=ref.step, rank=r, world=world, shared=shared) for r in range(world)]
        for it in range(1, 26):
            for s in ranks:
                s.phase_a(check=(it == 25))
            for s in ranks:
                s.phase_b()
        for s in ranks:
            s.wait_all_gather()
        beta = torch.cat([s.beta[:B, : s.rs.genes_owned] for s in ranks], 1)
        loss = sum(s.loss[:B] for s in ranks) / n + sum(s.penalty_owned() for s in ranks)
        vhl = [s.vhl.local.float() for s in ranks]
    assert beta.shape == (B, p) and (beta - ref.beta).abs().max() < 2e-4
    assert all(torch.equal(vhl[0], v) for v in vhl[1:])              # every rank holds the same extrapolated point
    # the loss of iteration 25 is taken at its extrapolated point v, the comparator reports the objective at the same iteration
    assert (loss - ref.objective).abs().max() < 5e-3


def test_batched_solver_accepts_prepared_operands():
    from consensusml_b200.models import lasso_large as LL
    x, y, lams, W = _batched_problem(96, 130, 5, 4)
    with _native.emulation(seed=2):
        a = LL.BatchedLassoLarge(x, y, lams, weights=W, step=0.5)
        b = LL.BatchedLassoLarge(None, y, lams, weights=W, step=0.5, n_total=96, prepared=(a.xb, a.xt))
        ra, rb = a.run(6, 0.0, 3), b.run(6, 0.0, 3)
    assert torch.equal(ra.beta, rb.beta)
    with pytest.raises(ValueError):
        LL.BatchedLassoLarge(None, y, lams, prepared=(a.xb, a.xt), n_total=96)


def test_random_ragged_shapes():
    """A seeded sample of awkward shapes (single rows / columns, sizes one off the tile edges, up to 8 owners, B beyond one
    n-tile) through both GEMM kernels and the fused reduce-scatter; 150 such draws passed when this was written."""
    import random
    rnd = random.Random(7)
    for _ in range(10):
        M = rnd.choice([1, 7, 127, 128, 129, 255, 257, 300]); N = rnd.choice([1, 5, 31, 33, 255, 257, 400, 769]); K = rnd.choice([64, 65, 72, 136, 200])
        variant, seed, sms = rnd.choice([1, 2]), rnd.randint(1, 50), rnd.choice([2, 4, 6])
        x, z = torch.randn(M, K), torch.randn(N, K)
        with _native.emulation(seed=seed, sms=sms):
            G.set_variant(variant)
            try:
                xa, na = G.to_bf16_rows(x)
                xb, nb = G.to_bf16_rows(z)
                d = G.gemm_tn(xa, [xb], N, epilogue=G.EPI_SQDIST, norm_a=na, norm_b=[nb])
            finally:
                G.set_variant(0)
        ref = (na[:, None] + nb[None, :] - 2 * xa.float() @ xb.float().t()).clamp_min(0)
        assert (d - ref).abs().max() < 3e-4, (M, N, K, variant, seed, sms)
    for _ in range(5):
        world, B, p, K = rnd.choice([1, 3, 5, 8]), rnd.choice([32, 96, 160, 320]), rnd.choice([32, 100, 257, 600]), rnd.choice([64, 100, 520, 1100])
        use_tma, split, seed, sms = rnd.choice([0, 1]), rnd.choice([0, 1]), rnd.randint(1, 50), rnd.choice([2, 4, 8])
        outs, _, ref, _, gpr = _run_gemm_rs(p, B, K, world, use_tma, split, seed, sms)
        scale = float(ref.abs().max())
        for dd in range(world):
            lo, hi = min(p, dd * gpr), min(p, (dd + 1) * gpr)
            if hi > lo:
                assert (outs[dd][: hi - lo] - ref[lo:hi]).abs().max() < 3e-5 * scale, (p, B, K, world, use_tma, split, seed, sms, dd)
            if hi - lo < gpr:
                assert float(outs[dd][hi - lo:].abs().max()) == 0.0


def test_consumers_of_the_gemm_run_their_kernel_branch_under_emulation(monkeypatch):
    """Code that only a GPU used to reach: k-means for large cohorts (tcgen05 distances + assign / update kernels), the forest
    proximity as one-hot GEMMs, the MLP's two products, the large-n SVM's kernel columns."""
    from consensusml_b200.models import kmeans as KM
    from consensusml_b200.models import mlp as MLP
    from consensusml_b200.models import svm_large as SL
    g = torch.Generator().manual_seed(3)
    centers = torch.randn(3, 80, generator=g) * 4
    x = centers[torch.arange(150) % 3] + torch.randn(150, 80, generator=g)
    monkeypatch.setattr(KM, "_FUSED_LIMIT", 100)                    # force the large-cohort path on a small problem
    ref = KM.kmeans(x, 3, nstart=3, seed=5)                          # CPU path
    with _native.emulation(seed=2):
        _native.reset_calls()
        got = KM.kmeans(x, 3, nstart=3, seed=5)
        assert _native.calls.get("cml_gemm_tn_bf16", 0) >= 3 and _native.calls.get("cml_kmeans_assign", 0) >= 3
    assert abs(got.tot_withinss - ref.tot_withinss) < 2e-2 * ref.tot_withinss      # bf16 distances: same partition, slightly different sums
    perm_ok = len({(int(a), int(b)) for a, b in zip(got.labels, ref.labels)}) == 3
    assert perm_ok
    a = torch.randn(64, 96, generator=g).bfloat16()
    b = torch.randn(40, 96, generator=g).bfloat16()
    with _native.emulation(seed=4):
        _native.reset_calls()
        c = MLP._gemm_nt(a, b, 40)
        assert _native.calls.get("cml_gemm_tn_bf16") == 1
        k = SL.LargeSVC(kernel="radial")._columns_single(x, x[:70], 1.0 / 80)
    assert (c[:, :40] - a.float() @ b.float().t()).abs().max() < 1e-3
    assert (k - torch.exp(-torch.cdist(x.double(), x[:70].double()) ** 2 / 80).float()).abs().max() < 2e-2


def test_stability_selection_on_the_batched_solver():
    """Bootstrap x lambda lanes through the batched sharded solver: the informative genes are selected in (almost) every
    replicate, the noise genes are not; the kernel path agrees with the torch path lane for lane."""
    from consensusml_b200.models import lasso_large as LL
    x, y, _, _ = _batched_problem(160, 96, 4, 11)
    lmax = LL.lambda_max(x, y, 160)
    lams = [0.35 * lmax, 0.15 * lmax]
    ref = LL.stability_selection_large(x, y, lams, n_boot=10, seed=3, max_iter=60, tol=0.0, backend="unfused")
    assert ref["frequency"].shape == (2, 96) and ref["n_total"] == 160
    top = torch.argsort(ref["max_frequency"], descending=True)[:4]
    assert set(top.tolist()) <= set(range(8)) and float(ref["max_frequency"][top].min()) >= 0.8
    assert float(ref["max_frequency"][20:].mean()) < 0.35
    with _native.emulation(seed=6):
        got = LL.stability_selection_large(x, y, lams, n_boot=10, seed=3, max_iter=60, tol=0.0, backend="fused")
    assert (got["frequency"] - ref["frequency"]).abs().max() <= 0.1 + 1e-9       # a coefficient at the threshold may flip in one replicate
