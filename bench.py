#!/usr/bin/env python
"""Flagship benchmark: stability-selection logistic lasso (glmnet-semantics 100-lambda paths).

One step = one ``stability_selection`` call of the public API on a fresh cohort of the reference's
modelling shape (93 training samples x 1984 DEG genes, the matrix behind the reference's headline
lasso result; cml_targetaml_seanalysis.Rmd:94-101): REPS bootstrap replicates per GPU, each a full
pathwise coordinate-descent fit over the reference's grid ``10^seq(10,-2,length=100)``, followed by
the per-gene selection frequencies (all-reduced over ranks when N > 1).  The metric is complete
path fits per second, summed over all GPUs (BASELINE.json: "fits/sec 1/2/4/8 GPU"); weak scaling -
every GPU always fits REPS replicates per step.

Contract: ``python bench.py --gpus N --steps K --warmup W`` prints ONE JSON line on rank 0.
``--impl reference`` reports that the reference cannot be run (it has no build system and its
modelling script imports xgboost, which is not installed).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_TRAIN, P_GENES, N_LAMBDA = 93, 1984, 100
# 8880 = 3 x (148 SMs x 20 resident path-fit warps): three full waves of the path kernel on a B200 (8192, used for the
# numbers in profiles/ up to this change, left the third wave 77 % full)
REPS_PER_GPU = int(os.environ.get("CML_BENCH_REPS", 8880))
METRIC = "lasso_path_fits_per_sec"


def reference_arm():
    print(json.dumps({
        "impl": "reference",
        "unavailable": "reference is R notebooks + scripts/model_comp.py: no setup.py/pyproject (pip install fails), "
                       "R/glmnet/xgboost are not installed and model_comp.py imports xgboost at module top",
    }))
    return 0


def synth_cohort(seed, n=N_TRAIN, p=P_GENES):
    """A cohort shaped like the reference's training matrix: log2-CPM-like genes with a planted
    100-gene weak class signal (55 / 38 class balance as in the reference's training split)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    y = np.zeros(n, np.float32)
    y[: int(round(n * 55 / 93))] = 1
    rng.shuffle(y)
    mu = rng.normal(2.0, 2.5, p)
    sd = rng.uniform(0.6, 2.2, p)
    x = (mu + sd * rng.standard_normal((n, p))).astype(np.float32)
    sig = rng.choice(p, 100, replace=False)      # weak, diffuse signal: ~1000 CD passes and df ~ 60 at the
    x[:, sig] += (y[:, None] - 0.5) * rng.uniform(0.2, 0.7, 100).astype(np.float32) * sd[sig].astype(np.float32)  # grid end, like the real cohort
    return x, y


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reps", type=int, default=REPS_PER_GPU)
    ap.add_argument("--cpu-baseline-fits", type=int, default=16)
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm()
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch
    import torch.distributed as dist
    from consensusml_b200 import _native
    from consensusml_b200.models import lasso
    from consensusml_b200.ops import lasso as K
    from consensusml_b200.runtime.clocks import ClockSampler

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _native.cuda()
    lam = K.default_lambda_grid()
    reps = args.reps
    total = args.warmup + args.steps

    # a distinct cohort per step, staged in pinned host memory (the e2e arm copies it every step)
    pool = []
    for s in range(total):
        x, y = synth_cohort(s)          # every rank bootstraps the SAME cohort (its replicates differ by seed): that is how
                                        # a stability-selection job is spread over GPUs, and it keeps the ranks' work balanced
        pool.append((torch.from_numpy(x).pin_memory(), torch.from_numpy(y).pin_memory()))
    flush_buf = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    lib = _native.cuda()

    def flush():
        lib.cml_l2_flush(_native.ptr(flush_buf), flush_buf.numel(), _native.stream_ptr())

    def step_device(xg, yg, s):
        out = lasso.stability_selection(xg, yg, lam, n_boot=reps, seed=17 + s + 100003 * rank, batch=reps)
        freq = out["freq"]
        if world > 1:
            dist.all_reduce(freq)
            freq /= world
        return freq.amax(dim=(0, 1))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- arm A: device-timed (inputs already resident) ----------------
    dev_pool = [(x.to(dev), y.to(dev)) for x, y in pool]
    for s in range(args.warmup):
        step_device(*dev_pool[s], s)
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local_rank) as clk:
        barrier()
        for i in range(args.steps):
            flush()
            ev[i][0].record()
            res = step_device(*dev_pool[args.warmup + i], args.warmup + i)
            ev[i][1].record()
        barrier()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = world * reps * args.steps / (ms_total / 1e3)
    sel_genes = int((res > 0.6).sum().item())

    # ---------------- arm B: end to end through the public API ----------------
    h2d = d2h = 0
    barrier()
    e2e_s = 0.0
    for i in range(args.steps):
        flush()
        torch.cuda.synchronize()
        xp, yp = pool[args.warmup + i]
        t0 = time.perf_counter()
        xg = xp.to(dev, non_blocking=True)
        yg = yp.to(dev, non_blocking=True)
        host_res = step_device(xg, yg, args.warmup + i).cpu()      # device -> host read of the step's result
        e2e_s += time.perf_counter() - t0
        h2d = xp.numel() * xp.element_size() + yp.numel() * yp.element_size()
        d2h = host_res.numel() * host_res.element_size()
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * reps * args.steps / float(t.item())

    # ---------------- context: the float64 C++ restatement of glmnet on one CPU core ----------------
    cpu_fits_per_s = None
    if rank == 0 and args.cpu_baseline_fits > 0:
        x, y = pool[0]
        nb = args.cpu_baseline_fits
        from consensusml_b200.ops.replicate import replicate_weights
        w = replicate_weights(nb, N_TRAIN, seed=1).double()
        t0 = time.perf_counter()
        K.lognet_path(x.double(), y.double(), w, lam, want_eta=False)
        cpu_fits_per_s = nb / (time.perf_counter() - t0)

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "path_fits/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"model": "logistic lasso, glmnet-semantics pathwise CD, stability selection",
                       "n_samples": N_TRAIN, "n_genes": P_GENES, "n_lambda": N_LAMBDA,
                       "lambda_grid": "10^seq(10,-2,length=100)", "replicates_per_gpu_per_step": reps,
                       "global_batch": reps * world, "parallelism": f"replicate-parallel x{world} + freq all-reduce",
                       "l2": "flushed (256 MiB write) before every timed step; inputs < L2",
                       "selected_genes_last_step": sel_genes},
            "clocks": clk.summary(),
            "e2e": {"value": e2e_value, "unit": "path_fits/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": 2 * args.steps,
            "gpu_launches_note": "per step: cml_replicate_weights + cml_lognet_path (ours); torch zeros/transpose/amax excluded",
            "cpu_oracle_fits_per_s_1core": cpu_fits_per_s,
        }))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
