"""Probe: symmetric memory across ranks, peer pointers, signal pads, NVLS multicast (run under torchrun)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from consensusml_b200.parallel import dist as D, symm
rank, world, dev = D.init()
buf = symm.alloc((1024,), torch.float32, dev)
buf.local.fill_(float(rank + 1))
D.barrier()
vals = [float(buf.peer_view(r)[0]) for r in range(world)]
print(f"rank {rank}/{world} dev {dev} peer_ptrs {[hex(p) for p in buf.peer_ptrs]} signal {[hex(p) for p in buf.signal_ptrs][:2]} "
      f"multicast {hex(buf.multicast_ptr)} peer values {vals}", flush=True)
h = buf.handle
if h is not None:
    print(rank, "signal_pad_size", getattr(h, "signal_pad_size", None), "buffer_size", getattr(h, "buffer_size", None), flush=True)
D.barrier()
D.shutdown()
