"""Run under torchrun (gloo, 2+ ranks): sample-sharded normalisation and gene-sharded voom against the single-process
functions.  The per-gene / per-sample kernels run on the host execution model (CPU tensors inside _native.emulation())."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

from consensusml_b200 import _native
from consensusml_b200.io.synthetic import nb_counts
from consensusml_b200.parallel import dist as D
from consensusml_b200.parallel.dist import shard_bounds
from consensusml_b200.stats import de as DE
from consensusml_b200.stats import normalize as N
from consensusml_b200.stats import sharded as SH


REF_COHORT = None


def main():
    c, g = nb_counts(700, 22, seed=4, n_de=80)
    ct = torch.from_numpy(c)
    with _native.emulation():
        ref_n = N.tmm_normalize(ct, 1.0, 5)
        ref_d = DE.voom_de(ct, g)
        from consensusml_b200 import pipeline as _pl
        from consensusml_b200.runtime import config as _cfg
        global REF_COHORT
        REF_COHORT = _pl.cohort_from_counts(ct, [f"ENSG{i:011d}.1" for i in range(700)], [f"s{i}" for i in range(22)], g,
                                            np.arange(22) % 3 != 0, _cfg.Config())
    rank, world, dev = D.init("gloo")
    lo, hi = shard_bounds(22, world, rank)
    with _native.emulation():
        sn = SH.tmm_normalize_sharded(ct[:, lo:hi].contiguous(), 1.0, 5)
        sd = SH.voom_de_sharded(ct[:, lo:hi].contiguous(), g)
    ok = bool(torch.equal(sn["keep"], ref_n.keep))
    ok &= bool(torch.allclose(sn["lib_size"], ref_n.lib_size)) and bool(torch.allclose(sn["norm_factors"], ref_n.norm_factors, rtol=1e-9, atol=1e-12))
    ok &= bool(torch.allclose(sn["logcpm"], ref_n.logcpm[:, lo:hi], rtol=1e-5, atol=1e-5))
    ok &= bool(np.array_equal(sd.keep.numpy(), ref_d.keep.numpy())) and bool(np.allclose(sd.logFC, ref_d.logFC, rtol=1e-8, atol=1e-10))
    ok &= bool(np.allclose(sd.t, ref_d.t, rtol=1e-7, atol=1e-9)) and bool(np.allclose(sd.adj_P_Val, ref_d.adj_P_Val, rtol=1e-6, atol=1e-12))
    ok &= bool(np.array_equal(sd.deg_order, ref_d.deg_order)) and abs(sd.df_prior - ref_d.df_prior) < 1e-6 * ref_d.df_prior
    # the whole counts -> DEG cohort front half, sample-sharded, against the single-process function
    from consensusml_b200 import pipeline
    from consensusml_b200.runtime import config as CFG
    cfg = CFG.Config()
    ids = [f"ENSG{i:011d}.1" for i in range(700)]
    sids = [f"s{i}" for i in range(22)]
    train = np.arange(22) % 3 != 0
    with _native.emulation():
        co_s = pipeline.cohort_from_counts_sharded(ct[:, lo:hi].contiguous(), ids, sids, g, train, cfg)
    ok &= co_s.gene_ids == REF_COHORT.gene_ids and bool(torch.allclose(co_s.assay, REF_COHORT.assay, rtol=1e-5, atol=1e-5))
    ok &= co_s.n_samples == 22 and co_s.metadata["sharded_over"] == world
    t = torch.tensor([1.0 if ok else 0.0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("STATS DIST OK" if float(t) == 1.0 else "STATS DIST FAILED", "DEGs", len(sd.deg_order), flush=True)
    D.shutdown()
    sys.exit(0 if float(t) == 1.0 else 1)


if __name__ == "__main__":
    main()
