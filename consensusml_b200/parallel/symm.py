"""Symmetric (peer-mapped) device buffers over NVLink.

Every rank allocates the same buffer, the ranks exchange handles once, and afterwards each rank holds the
device address of every peer's copy.  The fused kernels take those addresses (``peer_ptrs``) plus the
addresses of a small per-rank signal pad and issue plain / atomic stores to them from inside the kernel -
that is the whole transport.  Built on ``torch.distributed._symmetric_memory`` (CUDA VMM + fabric handles).
With one rank the "peers" are the local buffer, so the same kernels run unchanged on a single GPU."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch
import torch.distributed as dist


@dataclass
class SymmBuffer:
    local: torch.Tensor            # this rank's copy
    peer_ptrs: List[int]           # device address of every rank's copy (index = rank)
    signal_ptrs: List[int]         # device address of every rank's signal pad
    multicast_ptr: int             # NVLS multicast address (0 if unavailable)
    rank: int
    world: int
    handle: object = None

    def barrier(self, channel: int = 0) -> None:
        if self.handle is not None:
            self.handle.barrier(channel=channel)
        elif dist.is_initialized():
            dist.barrier()

    def peer_view(self, r: int) -> torch.Tensor:
        """Tensor view of rank r's copy (for tests / un-fused fallbacks)."""
        if self.handle is None:
            views = getattr(self, "_peer_views", None)
            return views[r] if views is not None else self.local
        return self.handle.get_buffer(r, tuple(self.local.shape), self.local.dtype)


_SIGNAL_PAD_WORDS = 256


def poison_enabled() -> bool:
    """``CML_SYMM_POISON=1``: debug mode of the fused paths - buffers whose contents are dead between two epochs
    (the idle half of a double-buffered landing area, a staging buffer before its copiers run) are filled with NaN, so that a
    consumer that reads before its producer's flag / signal for THIS epoch has arrived returns NaN instead of last epoch's
    plausible numbers.  Accumulation targets cannot be poisoned (they must be zero); they are checked by their own tests."""
    import os
    return os.environ.get("CML_SYMM_POISON", "0") == "1"


def poison_(t: torch.Tensor) -> None:
    if t.dtype.is_floating_point:
        t.fill_(float("nan"))
    else:
        t.fill_(-1)


_SHM_SEQ = 0


def _alloc_host_shared(shape, dtype, world: int, rank: int) -> SymmBuffer:
    """Several PROCESSES on the host execution model (torchrun + gloo + ``_native.emulation()``): every rank backs its buffer
    and its signal pad with a POSIX shared-memory file and maps all of its peers' files, so the "peer pointers" the kernels get
    are real addresses of the other processes' buffers - the in-kernel stores, reductions and signals of the fused paths then
    cross process boundaries the way they cross NVLink.  (tests/dist_fused_check.py --emu.)"""
    import mmap
    import os
    global _SHM_SEQ
    _SHM_SEQ += 1
    token = [f"{os.getpid()}_{_SHM_SEQ}"]
    dist.broadcast_object_list(token, src=0)
    numel = 1
    for d in (shape if isinstance(shape, (tuple, list)) else (shape,)):
        numel *= int(d)
    esize = torch.empty(0, dtype=dtype).element_size()
    sizes = {"buf": max(4096, (numel * esize + 4095) // 4096 * 4096), "pad": 4096}
    import tempfile
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else tempfile.gettempdir()   # any shared mapping will do
    path = lambda kind, r: os.path.join(base, f"cml_symm_{token[0]}_{kind}_{r}")
    for kind, size in sizes.items():
        with open(path(kind, rank), "wb") as fh:
            fh.truncate(size)                                   # zero pages
    dist.barrier()
    maps, tens = {}, {}
    for kind, size in sizes.items():
        for r in range(world):
            fd = os.open(path(kind, r), os.O_RDWR)
            mm = mmap.mmap(fd, size)
            os.close(fd)
            maps[(kind, r)] = mm
            tens[(kind, r)] = torch.frombuffer(mm, dtype=torch.uint8, count=size)
    dist.barrier()
    for kind in sizes:
        os.unlink(path(kind, rank))                             # the mappings keep the memory alive
    view = lambda r: tens[("buf", r)][: numel * esize].view(dtype).view(shape)
    buf = SymmBuffer(view(rank), [tens[("buf", r)].data_ptr() for r in range(world)], [tens[("pad", r)].data_ptr() for r in range(world)], 0,
                     rank, world, None)
    buf._keep = (maps, tens)
    buf._peer_views = [view(r) for r in range(world)]
    buf._pad = tens[("pad", rank)].view(torch.int32)
    return buf


def alloc(shape, dtype, device) -> SymmBuffer:
    """Collective: every rank calls with the same shape / dtype."""
    from .. import _native
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    if world > 1 and torch.device(device).type != "cuda" and _native.emulating():
        return _alloc_host_shared(shape, dtype, world, rank)
    if world == 1 or torch.device(device).type != "cuda":
        t = torch.zeros(shape, dtype=dtype, device=device)
        pad = torch.zeros(_SIGNAL_PAD_WORDS, dtype=torch.int32, device=device)
        buf = SymmBuffer(t, [t.data_ptr()] * world, [pad.data_ptr()] * world, 0, rank, world, None)
        buf._pad = pad
        return buf
    import torch.distributed._symmetric_memory as sm
    t = sm.empty(shape, dtype=dtype, device=device)
    t.zero_()
    h = sm.rendezvous(t, dist.group.WORLD)
    mc = 0
    try:
        mc = int(h.multicast_ptr or 0)
    except Exception:
        mc = 0
    # Our kernels keep their completion words in a symmetric buffer of their own: torch's signal pad belongs to
    # torch (its barrier() lives there), and sharing it deadlocks.
    sig = sm.empty((_SIGNAL_PAD_WORDS,), dtype=torch.int32, device=device)
    sig.zero_()
    hs = sm.rendezvous(sig, dist.group.WORLD)
    torch.cuda.synchronize()
    h.barrier(channel=0)
    buf = SymmBuffer(t, [int(p) for p in h.buffer_ptrs], [int(p) for p in hs.buffer_ptrs], mc, rank, world, h)
    buf._pad, buf._pad_handle = sig, hs
    return buf


class LocalGroup:
    """``world`` ranks played inside ONE process: every "rank" gets a plain tensor and the addresses of all of them, which is
    all the fused kernels need.  For the host execution model (several emulated ranks on the CPU) and for single-GPU
    multi-rank tests; a real multi-GPU run uses ``alloc``."""

    def __init__(self, world: int, device="cpu"):
        self.world, self.device = world, device
        self.pads = [torch.zeros(_SIGNAL_PAD_WORDS, dtype=torch.int32, device=device) for _ in range(world)]

    def alloc(self, shape, dtype) -> List[SymmBuffer]:
        ts = [torch.zeros(shape, dtype=dtype, device=self.device) for _ in range(self.world)]
        out = []
        for r in range(self.world):
            buf = SymmBuffer(ts[r], [t.data_ptr() for t in ts], [q.data_ptr() for q in self.pads], 0, r, self.world, None)
            buf._keep = (ts, self.pads)
            out.append(buf)
        return out
