"""Sample SM clocks and throttle reasons while a timed region runs (B200_PROFILING.md 'clocks line')."""
from __future__ import annotations

import statistics
import threading
from typing import Dict, List

_REASONS = {
    0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
    0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
    0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
}


class ClockSampler:
    def __init__(self, device_index: int = 0, period_s: float = 0.05):
        self.idx = device_index
        self.period = period_s
        self.samples: List[int] = []
        self.reasons = set()
        self.power: List[float] = []
        self.max_mhz = None
        self._stop = threading.Event()
        self._thr = None
        self.error = None

    def _run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            while not self._stop.is_set():
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                try:
                    self.power.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
                except Exception:
                    pass
                try:
                    bits = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    bits = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for b, name in _REASONS.items():
                    if bits & b and name != "gpu_idle":
                        self.reasons.add(name)
                self._stop.wait(self.period)
        except Exception as exc:  # no NVML: report it instead of failing the bench
            self.error = repr(exc)

    def __enter__(self):
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._thr.join(timeout=2)

    def summary(self) -> Dict:
        out = {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
               "reasons": sorted(self.reasons), "samples": len(self.samples)}
        if self.power:
            out["power_w_max"] = max(self.power)
        if self.error:
            out["error"] = self.error
        return out


_GLOBAL: "ClockSampler | None" = None


def start_global(device_index: int = 0) -> "ClockSampler":
    """Start one sampler for the lifetime of a benchmark script; ``snapshot()`` then returns the record of the interval
    since the previous snapshot (so every printed measurement carries the clocks seen while IT ran)."""
    global _GLOBAL
    if _GLOBAL is None:
        _GLOBAL = ClockSampler(device_index)
        _GLOBAL.__enter__()
    return _GLOBAL


def snapshot() -> Dict:
    if _GLOBAL is None:
        return {}
    out = _GLOBAL.summary()
    _GLOBAL.samples.clear()
    _GLOBAL.reasons.clear()
    _GLOBAL.power.clear()
    return out
