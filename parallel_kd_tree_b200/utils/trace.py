"""Phase tracing (the reference only has one DEBUG wall-clock number, kdtree_sequential.cpp:186-191).

``KD_B200_TRACE=1`` (or ``trace.enable()``) makes the library record named ranges — CUDA events on the
current stream for device work, wall clock otherwise — and ``trace.report()`` prints them.  The CLI's
``--trace`` switch prints the table on stderr, so stdout stays the reference's protocol.
"""
from __future__ import annotations

import os
import sys
import time
from contextlib import contextmanager

import torch

_enabled = os.environ.get("KD_B200_TRACE", "0") not in ("", "0")
_ranges: list[tuple[str, object, object, float]] = []


def enable(on: bool = True) -> None:
    global _enabled
    _enabled = on


def enabled() -> bool:
    return _enabled


def reset() -> None:
    _ranges.clear()


@contextmanager
def span(name: str, device: torch.device | None = None):
    """Times the enclosed work; a no-op unless tracing is enabled."""
    if not _enabled:
        yield
        return
    cuda = device is not None and torch.device(device).type == "cuda"
    t0 = time.perf_counter()
    if cuda:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(torch.cuda.current_stream(device))
    try:
        yield
    finally:
        if cuda:
            b.record(torch.cuda.current_stream(device))
            _ranges.append((name, a, b, 0.0))
        else:
            _ranges.append((name, None, None, (time.perf_counter() - t0) * 1e3))


def rows() -> list[tuple[str, float, str]]:
    out = []
    for name, a, b, ms in _ranges:
        if a is not None:
            b.synchronize()
            out.append((name, a.elapsed_time(b), "device"))
        else:
            out.append((name, ms, "host"))
    return out


def report(file=None) -> None:
    file = file or sys.stderr
    for name, ms, kind in rows():
        print(f"[kd trace] {name:28s} {ms:10.3f} ms  ({kind})", file=file)
