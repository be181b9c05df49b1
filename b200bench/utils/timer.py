"""Step timers.

``host``  reproduces the reference's timer bit for bit: ``time.perf_counter`` around the step
          with no device synchronisation (``benchmarking/train_harness.py:361-385``).
``device`` (default) brackets every step with CUDA events on the compute stream and reads them
          after one synchronise at the end, then the harness takes the max over ranks.
"""
from __future__ import annotations

import time
from typing import List

import torch


class StepTimer:
    def __init__(self, mode: str, device: torch.device):
        self.mode = mode if device.type == "cuda" else "host"
        self.device = device
        self._host: List[float] = []
        self._events: List[tuple] = []
        self._t0 = 0.0
        self._e0 = None

    def start(self) -> None:
        if self.mode == "device":
            self._e0 = torch.cuda.Event(enable_timing=True)
            self._e0.record()
        self._t0 = time.perf_counter()

    def stop(self, keep: bool) -> float:
        dt = time.perf_counter() - self._t0
        if self.mode == "device":
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            if keep:
                self._events.append((self._e0, e1))
        if keep:
            self._host.append(dt)
        return dt

    def times(self) -> List[float]:
        """Seconds per kept step."""
        if self.mode == "device":
            torch.cuda.synchronize(self.device)
            return [a.elapsed_time(b) / 1e3 for a, b in self._events]
        return list(self._host)

    def host_times(self) -> List[float]:
        return list(self._host)
