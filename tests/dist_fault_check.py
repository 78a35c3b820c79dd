"""Fault injection for the signal protocol: rank 1 never pushes its part of a reduce-scatter.  Rank 0's wait
kernel must give up after CML_WAIT_TIMEOUT_S and stop the process (a trap - a CUDA error on a GPU) instead of spinning for
ever.  Run under torchrun with gloo on the host execution model."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from consensusml_b200 import _native
from consensusml_b200.parallel import dist as D
from consensusml_b200.parallel import fused as F


def main():
    os.environ["CML_WAIT_TIMEOUT_S"] = "2"
    with _native.emulation(seed=1, sms=4):
        rank, world, dev = D.init("gloo")
        n, p = 64, 96
        g = torch.Generator().manual_seed(rank)
        xb = F.to_bf16_padded(torch.randn(n, p, generator=g))
        R = torch.randn(n, 1, generator=g)
        rs = F.XtrReduceScatter(p, 1, dev)
        if rank == 0:
            rs.launch(xb, R)
            print("[0] pushed, waiting for a peer that will never push", flush=True)
            t0 = time.time()
            rs.wait()                                   # must trap
            print(f"[0] WAIT RETURNED after {time.time() - t0:.1f} s", flush=True)
        else:
            time.sleep(8)                               # alive, but silent
    sys.exit(0)


if __name__ == "__main__":
    main()
