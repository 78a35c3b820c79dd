"""Smoke run of bench/fused_paths.py::measure on the host execution model (torchrun + gloo, tiny shapes): every code path of
the bench - including the opt-in ones - executes and returns well-formed records.  The numbers mean nothing."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bench"))
import torch.distributed as dist

from consensusml_b200 import _native
from consensusml_b200.parallel import dist as D


def main():
    import faulthandler
    faulthandler.dump_traceback_later(900, exit=True)
    with _native.emulation(seed=int(os.environ.get("RANK", "0")) + 1, sms=8):
        rank, world, dev = D.init("gloo")
        import fused_paths as FP
        recs = FP.measure(rank, world, dev, rows_per_gpu=264, genes=136, gram_rows=96, reps=1, nodes=2, bins_=16, experimental=True)
        paths = [r["path"] for r in recs]
        need = {"xtr_reduce_scatter", "hist_allreduce", "hist_reduce_scatter_level", "gram_allgather_tcgen05_rbf", "xtr_gemm_reduce_scatter",
                "lasso_batched_iteration", "shard_transpose"}
        ok = need <= set(paths) and all(r["ms"] > 0 and r["n_gpus"] == world and "clocks" in r for r in recs)
        if rank == 0:
            print("paths:", paths)
            print("BENCH SMOKE OK" if ok else "BENCH SMOKE FAILED", flush=True)
        D.barrier()
        D.shutdown()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
