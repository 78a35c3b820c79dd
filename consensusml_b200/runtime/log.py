"""Structured run log, device timers / NVTX ranges and numerical failure checks.

The reference's only observability is printed scalars (log-losses scripts/model_comp.py:30-33, confusion matrices
cml_targetaml_seanalysis.Rmd:107-108) and two sklearn progress banners (the only wall-clock numbers it contains,
model_walkthrough.ipynb:195); its only failure handling is glmnet's ``jerr``, a ConvergenceWarning the author reacted to
by hand (model_walkthrough.ipynb:1456-1457) and an NA guard (Differential_Expression_and_Lasso.Rmd:321-324).
Here: one JSON object per line (stage, rank, device milliseconds, bytes, achieved GB/s), CUDA-event timing with an
NVTX range per stage (visible to ncu / nsys), and explicit finite-value / convergence-status checks."""
from __future__ import annotations

import json
import os
import time
from contextlib import contextmanager
from typing import Any, Dict, Optional

import torch


class RunLog:
    def __init__(self, path: Optional[str] = None, rank: int = 0, echo: bool = False):
        self.path, self.rank, self.echo = path, rank, echo
        self.records = []
        if path:          # every rank writes its own file (<path>.rank<r> for r > 0), possibly before rank 0 has created the directory
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)

    def event(self, kind: str, **fields) -> Dict[str, Any]:
        rec = {"t": time.time(), "rank": self.rank, "event": kind, **fields}
        self.records.append(rec)
        if self.path:
            with open(self.path if self.rank == 0 else f"{self.path}.rank{self.rank}", "a") as fh:
                fh.write(json.dumps(rec, default=str) + "\n")
        if self.echo and self.rank == 0:
            print(json.dumps(rec, default=str), flush=True)
        return rec

    @contextmanager
    def stage(self, name: str, bytes_moved: Optional[float] = None, flops: Optional[float] = None, **fields):
        """Times a stage on the device (CUDA events on the current stream, synchronised on both sides) inside an NVTX
        range, then logs milliseconds and, when given, achieved GB/s and TFLOP/s."""
        cuda = torch.cuda.is_available()
        if cuda:
            torch.cuda.synchronize()
            torch.cuda.nvtx.range_push(name)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
        t0 = time.perf_counter()
        ok = True
        try:
            yield
        except Exception as exc:
            ok = False
            self.event("stage_failed", stage=name, error=repr(exc), **fields)
            raise
        finally:
            if cuda:
                b.record()
                torch.cuda.synchronize()
                torch.cuda.nvtx.range_pop()
                ms = a.elapsed_time(b)
            else:
                ms = (time.perf_counter() - t0) * 1e3
            if ok:
                rec = {"stage": name, "device_ms": ms, "wall_ms": (time.perf_counter() - t0) * 1e3, **fields}
                if bytes_moved:
                    rec["gb_per_s"] = bytes_moved / ms / 1e6
                if flops:
                    rec["tflop_per_s"] = flops / ms / 1e9
                self.event("stage", **rec)


class NumericalFailure(RuntimeError):
    pass


def check_finite(name: str, *tensors: torch.Tensor) -> None:
    """Raise with the tensor's name when a NaN / Inf appears (checked after every outer iteration of the solvers)."""
    for i, t in enumerate(tensors):
        if t is not None and not bool(torch.isfinite(t).all()):
            bad = int((~torch.isfinite(t)).sum())
            raise NumericalFailure(f"{name}[{i}]: {bad} non-finite values out of {t.numel()}")


JERR_MEANING = {0: "ok", 8001: "all responses in one class (weighted)", 9001: "all responses in one class (weighted)",
                10000: "no positive observation weight"}


def describe_jerr(code: int) -> str:
    """glmnet-style status of a path fit (the framework's ``jerr`` equivalent, one per replicate)."""
    if code in JERR_MEANING:
        return JERR_MEANING[code]
    if code < -10000:
        return f"active-set capacity reached at lambda index {-code - 10000}; coefficients up to the previous lambda are valid"
    if code < 0:
        return f"no convergence within the pass limit at lambda index {-code}"
    return f"fatal error {code}"


def retry_replicates(fn, seeds, max_retries: int = 1):
    """Run ``fn(seed)`` per replicate, retrying failures: replicates are idempotent given their seed."""
    out, failed = {}, []
    for s in seeds:
        for attempt in range(max_retries + 1):
            try:
                out[s] = fn(s)
                break
            except NumericalFailure:
                if attempt == max_retries:
                    failed.append(s)
    return out, failed
