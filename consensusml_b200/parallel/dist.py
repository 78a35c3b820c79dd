"""Process-group plumbing: one process per GPU, ``torch.distributed`` (NCCL on GPUs, gloo on CPU).

The reference is single-process CPU code, so everything here is new.  NCCL is the control
plane (bootstrap, barriers) and the un-fused comparator; the data plane of the three fused paths is
symmetric memory + in-kernel peer stores (parallel/symm.py, parallel/fused.py)."""
from __future__ import annotations

import os
from contextlib import contextmanager
from typing import Optional

import torch
import torch.distributed as dist


def init(backend: Optional[str] = None, timeout_s: Optional[float] = None) -> tuple[int, int, torch.device]:
    """Initialise from the torchrun environment (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*).
    Returns (rank, world_size, device).  A single-process call needs no environment.

    ``timeout_s`` (default: ``CML_DIST_TIMEOUT_S`` or 600) is the collective watchdog: a rank that never arrives makes
    the others raise instead of hanging the node (the fused kernels add their own guard - tests run them under a
    ``faulthandler`` deadline because a spinning kernel cannot be interrupted by NCCL's watchdog)."""
    import datetime
    timeout = datetime.timedelta(seconds=float(timeout_s if timeout_s is not None else os.environ.get("CML_DIST_TIMEOUT_S", 600)))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and backend != "gloo"
    device = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if use_cuda:
            dist.init_process_group(backend or "nccl", device_id=device, timeout=timeout)
        else:
            dist.init_process_group(backend or "gloo", timeout=timeout)
    return rank, world, device


def world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def barrier() -> None:
    if dist.is_initialized():
        dist.barrier()
    if torch.cuda.is_available():
        torch.cuda.synchronize()


def all_reduce_(t: torch.Tensor, op=dist.ReduceOp.SUM) -> torch.Tensor:
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=op)
    return t


def max_over_ranks(value: float, device=None) -> float:
    """Device-timed numbers are reported as the maximum over ranks."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return value
    dev = device or (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu")
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def shard_bounds(n: int, world: int, r: int) -> tuple[int, int]:
    """Contiguous block partition of ``n`` items: rank r owns [lo, hi)."""
    base, rem = divmod(n, world)
    lo = r * base + min(r, rem)
    return lo, lo + base + (1 if r < rem else 0)


def equal_shard(n: int, world: int) -> int:
    """Rows per rank when every rank must hold the same padded count (symmetric buffers)."""
    return (n + world - 1) // world


@contextmanager
def device_timer(sync: bool = True):
    """CUDA-event timer on the current stream; ``elapsed_ms`` is filled on exit."""
    class _T:
        elapsed_ms = 0.0
    t = _T()
    if torch.cuda.is_available():
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if sync:
            torch.cuda.synchronize()
        a.record()
        yield t
        b.record()
        torch.cuda.synchronize()
        t.elapsed_ms = a.elapsed_time(b)
    else:
        import time
        t0 = time.perf_counter()
        yield t
        t.elapsed_ms = (time.perf_counter() - t0) * 1e3


def shutdown() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()
