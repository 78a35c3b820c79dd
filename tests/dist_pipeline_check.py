"""Run under torchrun (2+ ranks): pipeline.run spread over the ranks must reproduce the single-process result.
CPU / gloo by default; with --cuda every rank uses its own GPU over NCCL; with --emu gloo ranks run every stage on the kernel
entry points (host execution model)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

from consensusml_b200 import pipeline
from consensusml_b200.parallel import dist as D
from consensusml_b200.runtime import config as C


def main():
    out = sys.argv[1]
    cuda = "--cuda" in sys.argv
    if "--emu" in sys.argv:
        from consensusml_b200 import _native
        _native.emulation(seed=int(os.environ.get("RANK", "0")) + 1).__enter__()      # for the life of the process
    rank, world, dev = D.init(None if cuda else "gloo")
    ov = ["lasso.exclusion_reps=2", "lasso.stability_boot=6", "lasso.stability_chunk=2", "rf.n_trees=[90, 150]", "rf.chunk=40", "rf.proximity=true",
          'gbm.reps=[{"max_depth": 2, "nrounds": 2}, {"max_depth": 3, "nrounds": 3}]', "mlp.enabled=true", "mlp.epochs=20",
          "mlp.hidden=[8]", "mlp.lr=[0.003]", "resume=false", "checkpoint=false", "device=" + ("cuda" if cuda else "cpu")]
    cfg = C.load(None, ov + ["io.out_dir=" + os.path.join(out, "dist")])
    det = pipeline.run(cfg, return_details=True)
    sched = det["summary"]["schedule"]
    assert len(sched["rank_seconds"]) == world and all(s > 0 for s in sched["rank_seconds"]), sched   # every rank did some work
    # every rank holds the complete result
    objs = [None] * world
    dist.all_gather_object(objs, {k: np.asarray(v).tolist() for k, v in det["importances"].items()})
    assert all(o == objs[0] for o in objs)
    dist.barrier()
    if rank == 0:
        cfg1 = C.load(None, ov + ["io.out_dir=" + os.path.join(out, "single"), "dist.enabled=false"])
        ref = pipeline.run(cfg1, return_details=True)
        a, b = det["standtable"], ref["standtable"]
        assert list(a.columns) == list(b.columns)
        num = [c for c in a.columns if a[c].dtype.kind == "f"]
        tol = dict(rtol=1e-6, atol=1e-9) if "--emu" not in sys.argv else dict(rtol=1e-4, atol=1e-6)    # emulated thread order differs per seed
        assert np.allclose(a[num].to_numpy(float), b[num].to_numpy(float), equal_nan=True, **tol)
        for k in ("lasso", "svm1", "rf_90", "rf_150", "xgb1", "xgb2"):
            sa, sb = det["summary"][k], ref["summary"][k]
            for f in ("auc", "test_error", "oob_error"):
                if f in sa:
                    assert abs(sa[f] - sb[f]) < (1e-9 if "--emu" not in sys.argv else 1e-6), (k, f, sa[f], sb[f])
        assert det["summary"]["consensus"]["top"] == ref["summary"]["consensus"]["top"]
        assert det["summary"]["stability"] == ref["summary"]["stability"] and det["summary"]["stability"]["stable_genes"] > 0
        assert os.path.exists(os.path.join(out, "dist", "standouttable.csv"))
        print("PIPELINE DIST OK", sched["critical"], flush=True)
    dist.barrier()
    D.shutdown()


if __name__ == "__main__":
    main()
