"""The allocation-free path solver entry point, driven through the C ABI exactly as a foreign caller would.

``cml_lognet_path_static`` exists for callers that cannot create device memory at the moment they need a
fit (a plugin inside a host application that owns the GPU heap, a service with a fixed memory budget, a
loop that must not hit cudaMalloc's device-wide synchronisation).  Its work space is an 8 MiB array built
into the library, so the only memory it touches besides that array is what the caller passes in.  The test
allocates inputs and outputs itself (that is the caller's side), then checks that the static-scratch entry
point returns what the ordinary entry point returns - twice in a row with different data, because the
built-in array is shared between calls and still holds the first call's values during the second.

Collected last on purpose (file name sorts after every other test file): a stop-at-first-failure run has
already exercised everything else by the time it gets here.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

_I, _LL, _SZ, _F, _P = ctypes.c_int, ctypes.c_longlong, ctypes.c_size_t, ctypes.c_float, ctypes.c_void_p


def _lib():
    from consensusml_b200 import _build
    return _declare(ctypes.CDLL(str(_build.CUDA_LIB)))


def _declare(lib):
    lib.cml_last_error.restype = ctypes.c_char_p
    head = [_I] * 6 + [_P, _P, _LL, _P, _P, _LL, _P, _P, _F, _F, _I, _I, _I]
    lib.cml_lognet_path.argtypes = head + [_P, _SZ] + [_P] * 11 + [_P]
    lib.cml_lognet_path.restype = _I
    lib.cml_lognet_path_static.argtypes = head + [_P] * 11 + [_P]
    lib.cml_lognet_path_static.restype = _I
    lib.cml_lognet_scratch_bytes.argtypes = [_I, _I, _I]
    lib.cml_lognet_scratch_bytes.restype = _SZ
    lib.cml_lognet_static_scratch_bytes.restype = _SZ
    lib.cml_lognet_n_pad.argtypes = [_I]
    lib.cml_lognet_n_pad.restype = _I
    return lib


def _problem(seed, B, n, p, L):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, p, generator=g)
    beta = torch.zeros(p)
    beta[:5] = torch.tensor([1.2, -1.0, 0.8, -0.6, 0.5])
    y = (torch.rand(n, generator=g) < torch.sigmoid(x @ beta)).float()
    w = torch.ones(B, n)
    for b in range(B):
        w[b, torch.randperm(n, generator=g)[: n // 8]] = 0      # every replicate leaves a few samples out
    lam = torch.tensor(np.geomspace(0.3, 0.01, L), dtype=torch.float32)
    return x, y, w, lam


def _outputs(B, L, nx, n_pad, p, dev="cuda"):
    i32, f32 = dict(dtype=torch.int32, device=dev), dict(dtype=torch.float32, device=dev)
    return dict(m=torch.zeros(B, nx, **i32), ca=torch.zeros(B, L, nx, **f32), kin=torch.zeros(B, L, **i32),
                a0=torch.zeros(B, L, **f32), dev=torch.zeros(B, L, **f32), eta=torch.zeros(B, L, n_pad, **f32),
                freq=torch.zeros(L, p, **f32), nulldev=torch.zeros(B, **f32), npasses=torch.zeros(B, **i32),
                lmu=torch.zeros(B, **i32), jerr=torch.zeros(B, **i32))


def _run(lib, static, x, y, w, lam, nx=128, device="cuda"):
    """``device="cpu"`` is used by tests/test_emu_kernels.py, which drives the same two entry points of the emulated
    library (host pointers, no stream)."""
    B, n = w.shape
    p = x.shape[1]
    L = lam.numel()
    n_pad = lib.cml_lognet_n_pad(n)
    xt = torch.zeros(p, n_pad, device=device)
    xt[:, :n] = x.t().to(device)
    yp = torch.zeros(n_pad, device=device)
    yp[:n] = y.to(device)
    wp = torch.zeros(B, n_pad, device=device)
    wp[:, :n] = w.to(device)
    lam_d = lam.to(device)
    o = _outputs(B, L, nx, n_pad, p, device)
    ptr = lambda t: ctypes.c_void_p(t.data_ptr())
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream if device == "cuda" else 0)
    head = [B, n, n_pad, p, L, nx, ptr(xt), ptr(yp), 0, ptr(wp), None, 0, None, ptr(lam_d), 1.0, 1e-7, 0, 0, 100000]
    outs = [ptr(o[k]) for k in ("m", "ca", "kin", "a0", "dev", "eta", "freq", "nulldev", "npasses", "lmu", "jerr")]
    if static:
        rc = lib.cml_lognet_path_static(*head, *outs, stream)
    else:
        nbytes = lib.cml_lognet_scratch_bytes(B, n_pad, p)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=device)      # the ordinary path: caller-provided scratch
        rc = lib.cml_lognet_path(*head, ptr(scratch), nbytes, *outs, stream)
    assert rc == 0, lib.cml_last_error().decode()
    if device == "cuda":
        torch.cuda.synchronize()
    return o


def _dense(o, b, p):
    L, nx = o["ca"].shape[1], o["ca"].shape[2]
    out = torch.zeros(L, p)
    kin = o["kin"][b].cpu()
    idx = o["m"][b].cpu().long()
    ca = o["ca"][b].cpu()
    for l in range(L):
        k = int(kin[l])
        out[l, idx[:k]] = ca[l, :k]
    return out


def test_static_scratch_entry_matches_regular_entry_twice(monkeypatch):
    # the regular entry would give a batch of at most one fit per SM to the one-CTA-per-fit kernel (round 2); this test is about
    # the allocation-free entry, so both sides are pinned to the kernel they share (the two kernels are compared with each other
    # in tests/kernel_cases.py::case_lasso_cta_per_fit_*)
    monkeypatch.setenv("CML_LASSO_TEAM", "0")
    lib = _lib()
    assert lib.cml_lognet_static_scratch_bytes() == 8 << 20
    # p chosen so that fewer replicates fit in the built-in array than the batch holds: 8 MiB / (12 * 3000) = 233 < 300,
    # which exercises the "same warps take further replicates in turn" path as well
    for seed, (B, n, p, L) in enumerate([(300, 70, 3000, 12), (40, 120, 500, 16)]):
        x, y, w, lam = _problem(100 + seed, B, n, p, L)
        ref = _run(lib, False, x, y, w, lam)
        got = _run(lib, True, x, y, w, lam)
        assert int(ref["jerr"].abs().max()) == 0 and int(got["jerr"].abs().max()) == 0
        # Both entry points run the same kernel on the same inputs; only the address of the work slab and the number
        # of warps sharing the batch differ, and neither enters the arithmetic of a fit: the results are identical.
        # (Round 1 needed tolerances here because the kernel read one shared-memory word before writing it; see
        # tests/test_emu_kernels.py::test_lasso_path_kernel_does_not_depend_on_stale_shared_memory_or_thread_order.)
        for k in ("kin", "npasses", "lmu", "a0", "dev", "eta", "freq", "nulldev"):
            assert torch.equal(ref[k], got[k]), k
        for b in (0, B // 2, B - 1):
            assert torch.equal(_dense(ref, b, p), _dense(got, b, p))
        assert float(got["freq"].sum()) > 0


def test_static_scratch_entry_rejects_oversized_problem():
    lib = _lib()
    dummy = torch.zeros(8, device="cuda")
    ptr = ctypes.c_void_p(dummy.data_ptr())
    p_too_big = (8 << 20) // 12 + 1
    head = [1, 32, 32, p_too_big, 1, 8, ptr, ptr, 0, ptr, None, 0, None, ptr, 1.0, 1e-7, 0, 0, 10]
    rc = lib.cml_lognet_path_static(*head, *([ptr] * 11), None)
    assert rc == 7                                            # CML_E2BIG
    msg = lib.cml_last_error().decode()
    assert "built-in array" in msg and str(p_too_big) in msg
