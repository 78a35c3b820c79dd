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
_emu: Optional[ctypes.CDLL] = None
_emu_depth = 1 if os.environ.get("CML_EMU") == "1" else 0


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


def emulating() -> bool:
    """True while kernels are served by the host execution model (``libcml_emu.so``, csrc/emu/cuda_emu.h)."""
    return _emu_depth > 0


def on_device(t) -> bool:
    """Should this tensor be handled by the kernel library?  CUDA tensors always; CPU tensors only inside
    ``emulation()`` (then the same kernel source runs on the host, thread by thread)."""
    return bool(t.is_cuda) or _emu_depth > 0


class emulation:
    """Context manager: route the op wrappers to the emulated kernel library for CPU tensors.

    ``seed`` (or ``CML_EMU_SEED``) != 0 shuffles the order in which emulated threads run between synchronisation
    points and changes the garbage dynamic shared memory starts with; results that depend on it expose a race or an
    uninitialised read.  ``sms`` sets the multiprocessor count the launchers see (grid sizing)."""

    def __init__(self, seed: Optional[int] = None, sms: Optional[int] = None):
        # CML_EMU_SEED_SHIFT moves every seed a test passes: a sweep over it re-runs the suite under other thread orders,
        # other shared-memory garbage and other completion orders of the asynchronous units (tools/emu_seed_sweep.sh)
        shift = int(os.environ.get("CML_EMU_SEED_SHIFT", "0"))
        if shift:
            seed = (seed or 0) + shift
        self.env = {"CML_EMU_SEED": seed, "CML_EMU_SMS": sms}
        self.saved = {}

    def __enter__(self):
        global _emu_depth
        if os.environ.get("CML_EMU_POISON_EMPTY") == "1" and _emu_depth == 0:
            # fresh host pages are zero, device memory from torch.empty is not: fill every "empty" tensor with NaN / large integers
            # so that a kernel (or its wrapper) relying on zero-initialised scratch fails here as it would on a GPU
            import torch
            self._orig_empty = (torch.empty, torch.empty_like)

            def _fill(t):
                if t.numel():
                    if t.dtype.is_floating_point or t.dtype.is_complex:
                        t.fill_(float("nan"))
                    elif t.dtype == torch.bool:
                        t.fill_(True)
                    else:
                        t.fill_(torch.iinfo(t.dtype).max // 3)
                return t
            oe, oel = self._orig_empty
            torch.empty = lambda *a, **k: _fill(oe(*a, **k))
            torch.empty_like = lambda *a, **k: _fill(oel(*a, **k))
        if int(os.environ.get("WORLD_SIZE", "1")) > 1:
            # several emulated ranks as processes (torchrun): a kernel that waits for a peer's signal needs a real clock
            os.environ.setdefault("CML_EMU_REALTIME", "1")
        for k, v in self.env.items():
            if v is not None:
                self.saved[k] = os.environ.get(k)
                os.environ[k] = str(v)
        _emu_depth += 1
        return self

    def __exit__(self, *a):
        global _emu_depth
        _emu_depth -= 1
        if getattr(self, "_orig_empty", None) is not None:
            import torch
            torch.empty, torch.empty_like = self._orig_empty
            self._orig_empty = None
        for k, old in self.saved.items():
            if old is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = old


def emu() -> ctypes.CDLL:
    global _emu
    with _lock:
        if _emu is None:
            path = _build.build_emu()
            lib = ctypes.CDLL(str(path))
            lib.cml_last_error.restype = ctypes.c_char_p
            _declare(lib)
            _emu = lib
        return _emu


def cuda() -> ctypes.CDLL:
    """Load the sm_100a kernel library (building it first if absent and nvcc exists).  Inside ``emulation()`` this is
    the emulated library instead: same entry points (minus the tcgen05 GEMMs), host pointers, no streams."""
    global _cuda
    if _emu_depth > 0:
        return emu()
    with _lock:
        if _cuda is None:
            if not _build.CUDA_LIB.exists():
                try:
                    _build.build_cuda()
                except Exception as exc:  # pragma: no cover - build box without nvcc
                    raise NativeError(
                        f"libcml_b200.so is missing and could not be built ({exc}); "
                        "run `python -m consensusml_b200._build`") from exc
            # CML_B200_LIB: load an alternative build of the kernel library (tuning experiments)
            lib = ctypes.CDLL(os.environ.get("CML_B200_LIB") or str(_build.CUDA_LIB))
            lib.cml_last_error.restype = ctypes.c_char_p
            _declare(lib)
            _cuda = lib
        return _cuda


_I, _LL, _SZ, _F, _D, _P, _ULL = (ctypes.c_int, ctypes.c_longlong, ctypes.c_size_t, ctypes.c_float, ctypes.c_double,
                                 ctypes.c_void_p, ctypes.c_ulonglong)

# Explicit signatures: without them ctypes passes every Python int as a 32-bit C int, which corrupts
# size_t / long long / pointer-sized arguments.
_SIGNATURES = {
    "cml_l2_flush": ([_P, _SZ, _P], _I),
    "cml_device_info": ([_P, _P, _P, _P], _I),
    "cml_lognet_n_pad": ([_I], _I),
    "cml_lognet_grid": ([_I, _I], _I),
    "cml_lognet_scratch_bytes": ([_I, _I, _I], _SZ),
    "cml_lognet_path": ([_I] * 6 + [_P, _P, _LL, _P, _P, _LL, _P, _P, _F, _F, _I, _I, _I, _P, _SZ] + [_P] * 11 + [_P], _I),
    "cml_replicate_weights": ([_P, _I, _I, _I, _ULL, _I, _I, _I, _P], _I),
}


def _declare(lib) -> None:
    for name, (args, res) in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.argtypes = args
        fn.restype = res


def register_signature(name: str, args, res=ctypes.c_int) -> None:
    """Used by op modules to declare the C signature of the entry points they bind."""
    _SIGNATURES[name] = (list(args), res)
    for lib in (_cuda, _emu):
        if lib is not None:
            _declare(lib)


# Every op wrapper reports the return code of its entry point through check(); counting there gives the number of
# kernel-launching calls of OUR library per entry point (bench.py reports it as gpu_launches).
calls: dict = {}


def reset_calls() -> None:
    calls.clear()


def check(code: int, what: str) -> None:
    with _lock:                                    # tasks may run on several host threads (parallel/tasks.py)
        calls[what] = calls.get(what, 0) + 1
    if code != 0:
        msg = cuda().cml_last_error().decode(errors="replace")
        raise NativeError(f"{what} failed with code {code}: {msg}")


class _KeepAlivePtr(ctypes.c_void_p):
    """A pointer argument that owns a reference to the tensor it points into.  ``ptr(expr)`` is often called on a temporary
    (``ptr(x.float().contiguous())``); without the reference the temporary dies as soon as ``ptr`` returns, its memory goes
    back to the allocator, and the NEXT temporary of the same argument list can be handed the same block - before the launch
    that was supposed to read it is even enqueued.  Holding the tensor until the foreign call returns makes the free
    stream-ordered after the launch, which is what the caching allocator needs."""
    _keep = None


def ptr(t) -> ctypes.c_void_p:
    """Device/host pointer of a contiguous torch tensor (``None`` -> NULL).  The returned object keeps ``t`` alive."""
    if t is None:
        return ctypes.c_void_p(0)
    if not t.is_contiguous():
        raise ValueError("native kernels need contiguous tensors")
    p = _KeepAlivePtr(t.data_ptr())
    p._keep = t
    return p


def sm_count() -> int:
    """Multiprocessors the launchers size their grids for (the emulated count inside ``emulation()``)."""
    if _emu_depth > 0:
        return max(1, int(os.environ.get("CML_EMU_SMS", "4")))
    import torch
    return torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count


def synchronize() -> None:
    """``torch.cuda.synchronize()`` on a GPU; nothing to wait for on the host execution model (launches run to completion)."""
    if _emu_depth == 0:
        import torch
        torch.cuda.synchronize()


def stream_ptr() -> ctypes.c_void_p:
    import torch
    if _emu_depth > 0:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def np_ptr(a) -> ctypes.c_void_p:
    if a is None:
        return ctypes.c_void_p(0)
    if not a.flags["C_CONTIGUOUS"]:
        raise ValueError("native host routines need C-contiguous arrays")
    p = _KeepAlivePtr(a.ctypes.data)
    p._keep = a
    return p
