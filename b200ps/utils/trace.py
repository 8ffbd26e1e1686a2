"""Tracing helpers: NVTX ranges and the per-step wall-clock line of the reference.

The reference's only instrumentation is ``time.time()`` deltas printed as ``AvgTime: ..ms`` every 100
batches (example.py:143,157,175-183).  Here:

* ``nvtx_range(name)`` — a context manager that opens an NVTX range when ``B200PS_NVTX=1`` (so that an
  Nsight Systems / ncu timeline shows ``pull``, ``forward_backward``, ``push``, ``step`` ...), and costs a
  single dict lookup otherwise.  The trainers wrap their step phases with it.
* ``StepTimer`` — host timer that reproduces the ``AvgTime`` statistic: average milliseconds per step since
  the last report.
* ``device_time_ms(fn, steps, warmup)`` — CUDA-event timing of ``steps`` calls of ``fn`` after ``warmup``
  untimed calls, synchronised on both sides (the protocol every number in BASELINE.md follows).
"""
from __future__ import annotations

import contextlib
import os
import time

_NVTX = os.environ.get("B200PS_NVTX", "0") not in ("", "0")


@contextlib.contextmanager
def nvtx_range(name: str):
    if not _NVTX:
        yield
        return
    import torch
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


def nvtx_enabled() -> bool:
    return _NVTX


class StepTimer:
    """``AvgTime`` of example.py:175-183: mean wall milliseconds per step since the previous ``report()``."""

    def __init__(self):
        self._t0 = time.time()
        self._n = 0

    def tick(self, n: int = 1) -> None:
        self._n += n

    def report(self) -> float:
        now = time.time()
        avg = (now - self._t0) * 1000.0 / max(1, self._n)
        self._t0, self._n = now, 0
        return avg


def device_time_ms(fn, steps: int, warmup: int = 3, device=None) -> float:
    """Milliseconds for ``steps`` calls of ``fn()`` measured with CUDA events on the current stream."""
    import torch
    for _ in range(max(0, warmup)):
        fn()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(device)
    return float(e0.elapsed_time(e1))
