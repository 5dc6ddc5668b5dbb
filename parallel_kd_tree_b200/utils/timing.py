"""Timers: CUDA-event device timing and the reference's DEBUG wall-clock line."""
from __future__ import annotations

import time

import torch


class CudaTimer:
    """CUDA events on the current stream; ``ms`` after a synchronize (device time, not wall clock)."""

    def __init__(self, device=None):
        self.device = device
        self.start = torch.cuda.Event(enable_timing=True)
        self.stop = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        self.start.record(torch.cuda.current_stream(self.device))
        return self

    def __exit__(self, *exc):
        self.stop.record(torch.cuda.current_stream(self.device))
        return False

    @property
    def ms(self) -> float:
        self.stop.synchronize()
        return self.start.elapsed_time(self.stop)


class PhaseTimer:
    """Wall-clock phases; ``elapsed_line`` is the reference's DEBUG output (kdtree_sequential.cpp:186-191)."""

    def __init__(self):
        self.t0 = time.perf_counter()
        self.phases: dict[str, float] = {}
        self._last = self.t0

    def mark(self, name: str) -> None:
        now = time.perf_counter()
        self.phases[name] = self.phases.get(name, 0.0) + (now - self._last)
        self._last = now

    def elapsed_line(self) -> str:
        return f"elapsed time {time.perf_counter() - self.t0:g} second"
