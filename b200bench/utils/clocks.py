"""SM clock / throttle-reason sampling during a timed region (B200 profiling recipe).

A background thread polls NVML (pynvml) every ``period`` seconds; ``summary()`` returns the median
SM clock under load, the max SM clock and the set of throttle reasons seen.  The reference has no
such facility (pynvml is installed by ``docker/Dockerfile:48`` but never imported).
"""
from __future__ import annotations

import statistics
import threading
import time
from typing import Any, Dict, List

_REASONS = {
    0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
    0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
    0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
}


class ClockSampler:
    def __init__(self, device_index: int = 0, period: float = 0.1):
        self.idx, self.period = device_index, period
        self._stop = threading.Event()
        self._thr = None
        self.sm: List[int] = []
        self.power: List[float] = []
        self.reason_bits = 0
        self.sm_max = None
        self.error = None

    def _run(self) -> None:
        try:
            import pynvml

            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            while not self._stop.is_set():
                self.sm.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                try:
                    self.power.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1e3)
                    self.reason_bits |= int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    try:
                        self.reason_bits |= int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                    except Exception:
                        pass
                time.sleep(self.period)
        except Exception as e:  # no NVML (CPU box): report it, never fail the run
            self.error = repr(e)

    def start(self) -> "ClockSampler":
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()
        return self

    def stop(self) -> Dict[str, Any]:
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=2.0)
        return self.summary()

    def summary(self) -> Dict[str, Any]:
        reasons = [name for bit, name in _REASONS.items() if self.reason_bits & bit and name != "gpu_idle"]
        out: Dict[str, Any] = {
            "sm_mhz": int(statistics.median(self.sm)) if self.sm else None,
            "sm_max_mhz": self.sm_max,
            "reasons": reasons,
            "samples": len(self.sm),
        }
        if self.power:
            out["power_w_max"] = round(max(self.power), 1)
        if self.error:
            out["error"] = self.error
        return out
