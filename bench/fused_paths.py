#!/usr/bin/env python
"""Roofline report for the three fused compute+collective paths (and the plain tcgen05 GEMM) on a synthetic cohort
sharded across the ranks: per-GPU shard fixed (weak scaling), device-timed with CUDA events, max over ranks,
L2 flushed between repetitions.  Prints one JSON line per measurement on rank 0.

    python bench/fused_paths.py                       # 1 GPU
    torchrun --nproc-per-node N bench/fused_paths.py  # N GPUs

Roofline denominators come from MEASURED_PEAKS.json (HBM copy bandwidth, cuBLAS bf16 throughput) and the
measured NVLink peer bandwidth of the profiling recipe (770 GB/s per direction)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from consensusml_b200 import _native
from consensusml_b200.parallel import dist as D
from consensusml_b200.parallel import fused as F

NVLINK_GBS = 770.0


def peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    except Exception:
        return 6650.0, 1590.0, "fallback"


class _Stamp:
    """CUDA event on a GPU; host clock on the host execution model (a smoke run of this file, not a measurement)."""

    def __init__(self):
        self.ev = None if _native.emulating() else torch.cuda.Event(enable_timing=True)
        self.t = 0.0

    def record(self):
        if self.ev is not None:
            self.ev.record()
        else:
            import time
            self.t = time.perf_counter()

    def elapsed_time(self, other):
        return self.ev.elapsed_time(other.ev) if self.ev is not None else (other.t - self.t) * 1e3


def _empty_cache():
    if not _native.emulating():
        torch.cuda.empty_cache()


def timed(fn, reps, flush):
    ms = []
    for _ in range(reps):
        flush()
        D.barrier()
        a, b = _Stamp(), _Stamp()
        a.record()
        fn()
        b.record()
        _native.synchronize()
        ms.append(a.elapsed_time(b))
    ms.sort()
    return D.max_over_ranks(ms[len(ms) // 2])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=125000, help="samples per GPU")
    ap.add_argument("--genes", type=int, default=20000)
    ap.add_argument("--gram-rows", type=int, default=8192, help="samples per GPU for the Gram path")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--nodes", type=int, default=4)
    ap.add_argument("--bins", type=int, default=256)
    ap.add_argument("--skip", default="")
    ap.add_argument("--experimental", action="store_true", help="also time the kernels / modes that are still opt-in")
    args = ap.parse_args()
    rank, world, dev = D.init()
    measure(rank, world, dev, rows_per_gpu=args.rows, genes=args.genes, gram_rows=args.gram_rows, reps=args.reps, nodes=args.nodes,
            bins_=args.bins, skip=args.skip, experimental=args.experimental, echo=True)
    D.barrier()
    D.shutdown()


def measure(rank, world, dev, rows_per_gpu=125000, genes=20000, gram_rows=8192, reps=5, nodes=4, bins_=256, skip="", experimental=False,
            echo=False):
    """Time every fused path on a per-GPU shard of the given shape.  Returns the list of records (one dict per measurement,
    each with the SM clock / throttle record sampled WHILE it was being timed); with ``echo`` rank 0 also prints them."""
    import types
    from consensusml_b200.runtime.clocks import ClockSampler
    args = types.SimpleNamespace(rows=rows_per_gpu, genes=genes, gram_rows=gram_rows, reps=reps, nodes=nodes, bins=bins_, skip=skip,
                                 experimental=experimental)
    hbm, tf, how = peaks()
    lib = _native.cuda()
    fb = torch.empty((256 << 20) if not _native.emulating() else 4096, dtype=torch.uint8, device=dev)
    flush = lambda: lib.cml_l2_flush(_native.ptr(fb), fb.numel(), _native.stream_ptr())
    n, p = args.rows, args.genes
    out = []
    sampler = ClockSampler(dev.index or 0)
    sampler.__enter__()

    def emit(**kw):
        kw.update(n_gpus=world, peaks=how, clocks=sampler.summary())
        sampler.samples.clear()
        sampler.reasons.clear()
        if rank == 0 and echo:
            print(json.dumps(kw), flush=True)
        out.append(kw)

    g = torch.Generator(device="cpu").manual_seed(rank)
    # ------------------------------------------------------------------ path 2: X^T r + reduce-scatter
    if "xtr" not in args.skip:
        ld = (p + 7) // 8 * 8
        xb = torch.empty(n, ld, dtype=torch.bfloat16, device=dev).normal_()
        for nb in (1, 4):
            R = torch.randn(n, nb, device=dev)
            rs = F.XtrReduceScatter(p, nb, dev)

            def run():
                rs.launch(xb, R)
                rs.wait()
            run(); _native.synchronize(); rs.grad.barrier(); rs.grad.local.zero_(); _native.synchronize(); rs.grad.barrier()
            ms = timed(run, args.reps, flush)
            bytes_hbm = n * ld * 2 + n * nb * 4
            link_bytes = p * nb * 4 * (world - 1) / world
            t_roof = max(bytes_hbm / (hbm * 1e9), link_bytes / (NVLINK_GBS * 1e9), 2.0 * n * p * nb / (tf * 1e12))
            emit(path="xtr_reduce_scatter", nb=nb, rows_per_gpu=n, genes=p, ms=ms, hbm_gbs=bytes_hbm / ms / 1e6,
                 frac_of_hbm_roofline=t_roof / (ms / 1e3), nvlink_bytes=link_bytes)
            if world > 1:
                def run_un():
                    full = torch.matmul(xb.t()[:p].float(), R)
                    o = torch.empty(rs.genes_per_rank, nb, device=dev)
                    pad = torch.zeros(rs.genes_per_rank * world, nb, device=dev)
                    pad[:p] = full
                    dist.reduce_scatter_tensor(o, pad)
                # the un-fused comparator materialises an fp32 copy of X; time only on small shards
                if n * p <= 2e8:
                    emit(path="xtr_unfused_torch_nccl", nb=nb, ms=timed(run_un, args.reps, flush))
            del rs, R
        del xb
        _empty_cache()
    # ------------------------------------------------------------------ path 1: histogram + all-reduce
    if "hist" not in args.skip:
        n_pad = (n + 3) // 4 * 4
        bins = torch.randint(0, args.bins, (p, n_pad), dtype=torch.uint8, device=dev)
        node = torch.randint(0, args.nodes, (n,), device=dev).to(torch.int16)
        gh = torch.randn(n, 2, device=dev)
        H = F.HistAllReduce(p, args.bins, args.nodes, dev)
        for mc in ([True, False] if world > 1 else [False]):
            def run():
                H.build(bins, n, node, gh, args.nodes, use_multicast=mc)
            run(); H.clear(args.nodes)
            ms_list = []
            for _ in range(args.reps):
                flush(); D.barrier()
                a, b = _Stamp(), _Stamp()
                a.record(); run(); b.record(); _native.synchronize()
                ms_list.append(a.elapsed_time(b)); H.clear(args.nodes)
            ms_list.sort()
            ms = D.max_over_ranks(ms_list[len(ms_list) // 2])
            bytes_hbm = p * n_pad + n * 10
            link = args.nodes * p * args.bins * 8 * (world - 1) / world
            t_roof = max(bytes_hbm / (hbm * 1e9), link / (NVLINK_GBS * 1e9))
            emit(path="hist_allreduce", multicast=mc, rows_per_gpu=n, genes=p, bins=args.bins, nodes=args.nodes, ms=ms,
                 frac_of_roofline=t_roof / (ms / 1e3), nvlink_bytes=link, smem_atomics=2.0 * n * p, atomics_per_s=2.0 * n * p / ms * 1e3)
        del H
        _empty_cache()
        # the reduce-scatter-by-feature landing with sibling subtraction: a level with 2 * nodes open nodes histograms only the
        # `nodes` smaller children (half of the samples), lands 1 / world of the bytes per rank, then derives + scores on the owner
        E = F.HistEngine(p, args.bins, args.nodes, 2 * args.nodes, dev)
        node2 = torch.randint(0, 2 * args.nodes, (n,), device=dev).to(torch.int16)
        pair = torch.arange(args.nodes, device=dev, dtype=torch.int16)
        slot_of = torch.stack([pair, torch.full_like(pair, -1)], 1).flatten().contiguous()      # left children built, right derived
        count = torch.ones(2 * args.nodes, device=dev)
        E.levels[1].normal_()                                                                   # a parent level to subtract from

        def level():
            E.build(bins, n, node2, gh, slot_of, args.nodes, 4.0, 4.0)
            full = E.derive(1, 2 * args.nodes, slot_of, count)
            E.best_keys(full, 2 * args.nodes, 1.0, 0.0, 1.0)
        level()
        ms = timed(level, args.reps, flush)
        bytes_hbm = p * n_pad + n * 10
        link = args.nodes * p * args.bins * 8 * (world - 1) / world
        t_roof = max(bytes_hbm / (hbm * 1e9), link / (NVLINK_GBS * 1e9))
        emit(path="hist_reduce_scatter_level", rows_per_gpu=n, genes=p, bins=args.bins, open_nodes=2 * args.nodes, built_slots=args.nodes, ms=ms,
             frac_of_roofline=t_roof / (ms / 1e3), nvlink_bytes=link, smem_atomics=1.0 * n * p, atomics_per_s=1.0 * n * p / ms * 1e3,
             includes="build (half of the samples) + peer landing + wait + sibling derivation + split scoring of the owned features")
        del bins, E
        _empty_cache()
    # ------------------------------------------------------------------ path 3: all-gather -> tcgen05 Gram (RBF)
    if "gram" not in args.skip:
        m = args.gram_rows
        x = torch.randn(m, p, device=dev)
        cache = {}

        def run():
            F.gram_allgather(x, kernel="radial", n_total=m * world, cache=cache)
        run()
        # time only the GEMM (conversion + publication are set-up): re-issue it on the published blocks
        from consensusml_b200.ops import gemm as G
        xbuf, nrm = cache["xb"], cache["nrm"]
        ld = xbuf.local.shape[1]
        outbuf = torch.empty(m, (m * world + 3) // 4 * 4, device=dev)

        def gemm_only():
            if world > 1:
                stg = cache["stage"]
                G.gemm_tn_staged(xbuf.local, xbuf.peer_ptrs, m * world, m, ld, stg["stage"], stg["norm"], stg["flags"], stg.next_epoch(),
                                 rank, epilogue=G.EPI_RBF, alpha=1.0 / p, norm_a=nrm.local, peer_norm_ptrs=nrm.peer_ptrs, out=outbuf)
            else:
                G.gemm_tn(xbuf.local, None, m * world, epilogue=G.EPI_RBF, alpha=1.0 / p, norm_a=nrm.local, b_ptrs=xbuf.peer_ptrs,
                          norm_b_ptrs=nrm.peer_ptrs, rows_per_src=m, ldb=ld, out=outbuf)
        gemm_only()
        ms = timed(gemm_only, args.reps, flush)
        flops = 2.0 * m * m * world * ld
        link = (world - 1) * m * ld * 2
        t_roof = max(flops / (tf * 1e12), link / (NVLINK_GBS * 1e9))
        emit(path="gram_allgather_tcgen05_rbf", rows_per_gpu=m, genes=p, ms=ms, tflops=flops / ms / 1e9,
             frac_of_roofline=t_roof / (ms / 1e3), nvlink_bytes=link)
    if args.experimental:
        # ---- opt-in: tcgen05 GEMM fused with the reduce-scatter (replicate-batched gradient), B = 256 residual columns
        n, p, B = args.rows, args.genes, 256
        ld = (n + 7) // 8 * 8
        xt = torch.randn(p, ld, device=dev, dtype=torch.bfloat16)
        rt = torch.randn(B, ld, device=dev, dtype=torch.bfloat16)
        for use_tma in (False, True):
            op = F.XtrGemmReduceScatter(p, B, dev, use_tma_reduce=use_tma)

            def run():
                op.launch(xt, rt)
                op.wait()
            run()
            ms = timed(run, args.reps, flush)
            flops = 2.0 * p * B * ld
            byts = (p + B) * ld * 2.0
            link = 4.0 * p * B * (world - 1) / world
            t_roof = max(flops / (tf * 1e12), byts / (hbm * 1e9), link / (NVLINK_GBS * 1e9))
            emit(path="xtr_gemm_reduce_scatter", tma_reduce=use_tma, rows_per_gpu=n, genes=p, B=B, ms=ms, tflops=flops / ms / 1e9,
                 hbm_gbs=byts / ms / 1e6, frac_of_roofline=t_roof / (ms / 1e3), nvlink_bytes=link)
            del op
        # ---- the whole replicate-batched iteration built on it: eta product (tcgen05), residual split, GEMM -> reduce-scatter
        # gradient, proximal step + all-gather of the bf16 rows of v (models/lasso_large.BatchedLassoLarge), B = 256 lanes
        from consensusml_b200.models import lasso_large as LL
        ldp = (p + 7) // 8 * 8
        xb = torch.randn(n, ldp, device=dev, dtype=torch.bfloat16)
        S = LL.BatchedLassoLarge(None, torch.randint(0, 2, (n,), device=dev).float(), [1e-3] * B, n_total=n * world, step=1e-3,
                                 prepared=(xb, xt))

        def it():
            S.phase_a(False)
            S.phase_b()
        it()
        ms = timed(it, args.reps, flush)
        flops = 8.0 * B * n * p                                   # two products, each with hi / lo rows
        byts = 2.0 * (n * ldp + p * ld) + 2 * B * n * 4 * 2.0
        link = 4.0 * p * B * (world - 1) / world + 4.0 * B * p * (world - 1) / world
        t_roof = max(flops / (tf * 1e12), byts / (hbm * 1e9), link / (NVLINK_GBS * 1e9))
        emit(path="lasso_batched_iteration", rows_per_gpu=n, genes=p, lanes=B, ms=ms, tflops=flops / ms / 1e9,
             frac_of_roofline=t_roof / (ms / 1e3), nvlink_bytes=link, lane_iterations_per_s=B / ms * 1e3)
        del S, xb, xt, rt
        if world > 1 and "gram" not in args.skip:
            from consensusml_b200.ops import gemm as G
            m = args.gram_rows
            xg = torch.randn(m, p, device=dev)
            cache = {}
            for grouped in (False, True):
                G.set_staged_grouping(grouped)
                F.gram_allgather(xg, kernel="radial", n_total=m * world, cache=cache)
                ms = timed(lambda: F.gram_allgather(xg, kernel="radial", n_total=m * world, cache=cache), args.reps, flush)
                emit(path="gram_allgather_whole_call", staged_grouped=grouped, rows_per_gpu=m, genes=p, ms=ms)
            G.set_staged_grouping(False)
    if "transpose" not in args.skip:
        # ---- path 4: sample-shard -> gene-shard transpose, 32768 samples per GPU x 20k genes fp32 (2.6 GB per GPU)
        from consensusml_b200.parallel import transpose as TR
        m, p = (32768 if not _native.emulating() else 64), args.genes
        x = torch.randn(m, p, device=dev)
        plan = TR.ShardTranspose(m * world, p, dev)
        plan(x)
        ms = timed(lambda: plan(x), args.reps, flush)
        ms_nccl = timed(lambda: TR.shard_transpose_unfused(x, m * world), args.reps, flush)
        byts = 2.0 * m * p * 4                                  # read once + write once per GPU
        link = m * p * 4 * (world - 1) / world
        t_roof = max(byts / (hbm * 1e9), link / (NVLINK_GBS * 1e9))
        emit(path="shard_transpose", rows_per_gpu=m, genes=p, ms=ms, nccl_all_to_all_plus_transpose_ms=ms_nccl,
             gbs=byts / ms / 1e6, frac_of_roofline=t_roof / (ms / 1e3), nvlink_bytes=link)
    sampler.__exit__(None, None, None)
    return out


if __name__ == "__main__":
    main()
