#!/usr/bin/env python
"""Where does a lasso path fit spend its time?  Runs the tracing build of the path kernel (clock64 phase timers,
``cml_lognet_path_profile``) on the bench cohort at three occupancies and prints mean cycles per phase per fit."""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench
from consensusml_b200 import _native
from consensusml_b200.ops import lasso as K
from consensusml_b200.ops.replicate import replicate_weights

_I, _LL, _F, _P, _SZ = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p, ctypes.c_size_t
_native.register_signature("cml_lognet_path_profile", [_I, _I, _I, _I, _I, _I, _P, _P, _LL, _P, _P, _F, _F, _I, _I, _I, _P, _SZ,
                                                        _P, _P, _P, _P, _P, _P, _P, _P, _I, _P])
PHASES = ["setup", "strong_rule", "xv_refresh", "screened_sweeps", "active_sweeps", "irls_refresh", "kkt", "record"]


def main():
    lib = _native.cuda()
    dev = torch.device("cuda")
    x, y = bench.synth_cohort(0)
    x, y = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    n, p = x.shape
    lam = torch.tensor(K.default_lambda_grid(), dtype=torch.float32, device=dev)
    n_pad = lib.cml_lognet_n_pad(n)
    xt = torch.zeros(p, n_pad, device=dev)
    xt[:, :n] = x.t()
    yp = torch.zeros(n_pad, device=dev)
    yp[:n] = y
    P = _native.ptr
    for B, limit in [(148, 148), (148 * 8, 148 * 8), (8192, 0)]:
        w = replicate_weights(B, n, seed=3, device=dev, n_pad=n_pad)
        grid = lib.cml_lognet_grid(B, n_pad)
        scratch = torch.empty(lib.cml_lognet_scratch_bytes(B, n_pad, p) // 4, device=dev)
        kin = torch.zeros(B, len(lam), dtype=torch.int32, device=dev)
        a0 = torch.zeros(B, len(lam), device=dev)
        devr = torch.zeros(B, len(lam), device=dev)
        nd = torch.zeros(B, device=dev)
        npass = torch.zeros(B, dtype=torch.int32, device=dev)
        lmu = torch.zeros(B, dtype=torch.int32, device=dev)
        jerr = torch.zeros(B, dtype=torch.int32, device=dev)
        prof = torch.zeros(B, 8, dtype=torch.int64, device=dev)
        for it in range(2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = lib.cml_lognet_path_profile(B, n, n_pad, p, len(lam), 256, P(xt), P(yp), 0, P(w), P(lam), 1.0, 1e-7, 0, 0, 100000,
                                             P(scratch), scratch.numel() * 4, P(kin), P(a0), P(devr), P(nd), P(npass), P(lmu),
                                             P(jerr), P(prof), limit, _native.stream_ptr())
            _native.check(rc, "cml_lognet_path_profile")
            e1.record()
            torch.cuda.synchronize()
        cyc = prof.double().mean(0).cpu().numpy()
        print(json.dumps({"B": B, "warps_per_sm": min(B, grid if not limit else limit) / 148, "ms": e0.elapsed_time(e1),
                          "jerr_nonzero": int((jerr != 0).sum()), "mean_npasses": float(npass.float().mean()),
                          "Mcycles_per_fit": {k: round(v / 1e6, 2) for k, v in zip(PHASES, cyc)}, "total_Mcycles": round(cyc.sum() / 1e6, 1)}),
              flush=True)


if __name__ == "__main__":
    main()
