"""ctypes bindings to the two in-tree native libraries.

``cuda()`` returns ``libcml_b200.so`` (sm_100a kernels behind a C ABI, see ``include/cml_b200.h``);
``host()`` returns ``libcml_host.so`` (C++ float64 oracles / CPU baseline / parsers).

Policy: a tensor on a CUDA device is always served by the CUDA library.  If the library is
missing on a machine that has a GPU we raise - silently falling back to eager PyTorch would make
GPU tests pass without exercising any native code.
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional

from . import _build

_lock = threading.Lock()
_cuda: Optional[ctypes.CDLL] = None
_host: Optional[ctypes.CDLL] = None


class NativeError(RuntimeError):
    pass


def host() -> ctypes.CDLL:
    global _host
    with _lock:
        if _host is None:
            if not _build.HOST_LIB.exists() or os.environ.get("CML_REBUILD"):
                _build.build_host()
            _host = ctypes.CDLL(str(_build.HOST_LIB))
        return _host


def cuda_available() -> bool:
    import torch
    return torch.cuda.is_available()


def cuda() -> ctypes.CDLL:
    """Load the sm_100a kernel library (building it first if absent and nvcc exists)."""
    global _cuda
    with _lock:
        if _cuda is None:
            if not _build.CUDA_LIB.exists():
                try:
                    _build.build_cuda()
                except Exception as exc:  # pragma: no cover - build box without nvcc
                    raise NativeError(
                        f"libcml_b200.so is missing and could not be built ({exc}); "
                        "run `python -m consensusml_b200._build`") from exc
            lib = ctypes.CDLL(str(_build.CUDA_LIB))
            lib.cml_last_error.restype = ctypes.c_char_p
            _cuda = lib
        return _cuda


def check(code: int, what: str) -> None:
    if code != 0:
        msg = cuda().cml_last_error().decode(errors="replace")
        raise NativeError(f"{what} failed with code {code}: {msg}")


def ptr(t) -> ctypes.c_void_p:
    """Device/host pointer of a contiguous torch tensor (``None`` -> NULL)."""
    if t is None:
        return ctypes.c_void_p(0)
    if not t.is_contiguous():
        raise ValueError("native kernels need contiguous tensors")
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr() -> ctypes.c_void_p:
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def np_ptr(a) -> ctypes.c_void_p:
    if a is None:
        return ctypes.c_void_p(0)
    if not a.flags["C_CONTIGUOUS"]:
        raise ValueError("native host routines need C-contiguous arrays")
    return ctypes.c_void_p(a.ctypes.data)
