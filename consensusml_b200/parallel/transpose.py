"""Sample-shard <-> gene-shard transpose (the pipeline's all-to-all).

The per-gene statistics (voom / lmFit / eBayes, correlation ranks) want every sample of a block of genes; the model
kernels (lasso gradient, histograms, Gram tiles) want every gene of a block of samples.  Moving between the two
layouts is an all-to-all of sub-blocks plus a local transpose.  ``ShardTranspose`` does both in ONE kernel
(``cml_shard_transpose``, csrc/fused.cu path 4): tiles are transposed through shared memory and stored straight into
the owner rank's symmetric buffer over NVLink, then a per-source signal word is released; the receiver waits on the
signal words on its own stream.  ``shard_transpose_unfused`` is the NCCL / gloo ``all_to_all`` comparator and the CPU
path.

The reference has no distributed code (its only inter-process transport is CSV files between the
Python and R halves - david_Eda.ipynb:3059 -> Differential_Expression_and_Lasso.Rmd:105)."""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _native
from . import symm
from .dist import shard_bounds

_I, _LL, _P, _U = ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_uint
_native.register_signature("cml_shard_transpose", [_P, _LL, _I, _I, _P, _P, _LL, _I, _P, _I, _I, _P, _U, _I, _P])
_native.register_signature("cml_wait_signals", [_P, _I, _I, _U, _P])

SIG_TRANSPOSE = 24          # signal-pad words 24..31 (parallel/fused.py uses 0..23)


def _world_rank():
    if dist.is_initialized():
        return dist.get_world_size(), dist.get_rank()
    return 1, 0


class ShardTranspose:
    """Collective object: ``rows_total x cols`` matrix, row-sharded by ``shard_bounds(rows_total, world, r)``;
    ``__call__(x_local)`` returns this rank's column block transposed: ``[cols_owned, rows_total]``.

    Results alternate between two symmetric buffers, so an output stays valid until the call after the next one."""

    def __init__(self, rows_total: int, cols: int, device, dtype=torch.float32):
        assert torch.empty(0, dtype=dtype).element_size() == 4, "32-bit element types only"
        self.world, self.rank = _world_rank()
        self.rows_total, self.cols, self.dtype = rows_total, cols, dtype
        self.col_start = [shard_bounds(cols, self.world, r)[0] for r in range(self.world)] + [cols]
        self.row0, self.row1 = shard_bounds(rows_total, self.world, self.rank)
        self.cols_owned = self.col_start[self.rank + 1] - self.col_start[self.rank]
        max_cols = max(self.col_start[r + 1] - self.col_start[r] for r in range(self.world))
        self.ldy = (rows_total + 3) // 4 * 4
        self.out = [symm.alloc((max_cols, self.ldy), dtype, device) for _ in range(2)]
        self.done = torch.zeros(4, dtype=torch.int32, device=device)
        self._cs = (ctypes.c_int * (self.world + 1))(*self.col_start)
        self.epoch = 0

    def __call__(self, x_local: torch.Tensor) -> torch.Tensor:
        assert _native.on_device(x_local) and x_local.dtype == self.dtype and x_local.dim() == 2 and x_local.stride(1) == 1
        assert x_local.shape == (self.row1 - self.row0, self.cols), (tuple(x_local.shape), self.row1 - self.row0, self.cols)
        self.epoch += 1
        buf = self.out[self.epoch & 1]
        lib = _native.cuda()
        arr = lambda ptrs: (ctypes.c_void_p * len(ptrs))(*ptrs)
        rc = lib.cml_shard_transpose(_native.ptr(x_local), x_local.stride(0), x_local.shape[0], self.cols, self._cs,
                                     arr(buf.peer_ptrs), self.ldy, self.row0, arr(buf.signal_ptrs), self.rank, self.world,
                                     _native.ptr(self.done), self.epoch, SIG_TRANSPOSE, _native.stream_ptr())
        _native.check(rc, "cml_shard_transpose")
        rc = lib.cml_wait_signals(ctypes.c_void_p(buf.signal_ptrs[self.rank]), SIG_TRANSPOSE, self.world, self.epoch,
                                  _native.stream_ptr())
        _native.check(rc, "cml_wait_signals")
        return buf.local[: self.cols_owned, : self.rows_total]


def shard_transpose_unfused(x_local: torch.Tensor, rows_total: int) -> torch.Tensor:
    """Same result through ``all_to_all`` (NCCL / gloo) and a local transpose; also the single-process path."""
    world, rank = _world_rank()
    cols = x_local.shape[1]
    if world == 1:
        return x_local.t().contiguous()
    c0, c1 = shard_bounds(cols, world, rank)
    send = [x_local[:, slice(*shard_bounds(cols, world, r))].contiguous() for r in range(world)]
    recv = [torch.empty(shard_bounds(rows_total, world, r)[1] - shard_bounds(rows_total, world, r)[0], c1 - c0,
                        dtype=x_local.dtype, device=x_local.device) for r in range(world)]
    if x_local.is_cuda:
        dist.all_to_all(recv, send)
    else:                                                     # gloo: one gather per destination
        for r in range(world):
            dist.gather(send[r], recv if r == rank else None, dst=r)
    return torch.cat(recv, 0).t().contiguous()


def to_gene_shards(x_local: torch.Tensor, rows_total: int, plan: Optional[ShardTranspose] = None) -> torch.Tensor:
    """samples x genes (row block) -> genes x samples (gene block); fused on CUDA, all_to_all otherwise."""
    if (x_local.is_cuda and _native.cuda_available()) or _native.emulating():
        plan = plan or ShardTranspose(rows_total, x_local.shape[1], x_local.device, x_local.dtype)
        return plan(x_local)
    return shard_transpose_unfused(x_local, rows_total)
