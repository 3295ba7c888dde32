"""Tracing / profiling helpers.

The reference only has ``time.time()`` / ``MPI.Wtime()`` around whole runs
(``tests/smf_example/smf_grad_descent.py:105-109``, ``benchmark.py:44-46``).
Here timings are taken on the device with CUDA events and reduced with a max
over ranks; optional NVTX ranges make the steps visible to Nsight tools.
"""
from __future__ import annotations

import contextlib
import time
from typing import Callable, List

import torch

__all__ = ["device_time_ms", "StepTimer", "nvtx_range"]


def device_time_ms(fn: Callable[[], None], iters: int = 10, warmup: int = 3,
                   comm=None) -> float:
    """Average device time of ``fn`` in ms (CUDA events on the current
    stream, synchronise on both sides, max over the ranks of ``comm``)."""
    for _ in range(warmup):
        fn()
    if not torch.cuda.is_available():
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        ms = (time.perf_counter() - t0) * 1e3 / iters
    else:
        torch.cuda.synchronize()
        if comm is not None:
            comm.Barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
    if comm is not None and comm.size > 1:
        ms = max(comm.allgather(ms))
    return ms


class StepTimer:
    """Per-step device timer that never synchronises inside the loop: events
    are recorded around every step and read back at the end."""

    def __init__(self):
        self._pairs: List = []

    @contextlib.contextmanager
    def step(self):
        if torch.cuda.is_available():
            a = torch.cuda.Event(enable_timing=True)
            b = torch.cuda.Event(enable_timing=True)
            a.record()
            yield
            b.record()
            self._pairs.append((a, b))
        else:
            t0 = time.perf_counter()
            yield
            self._pairs.append((time.perf_counter() - t0) * 1e3)

    def times_ms(self) -> List[float]:
        if torch.cuda.is_available():
            torch.cuda.synchronize()
            return [a.elapsed_time(b) for a, b in self._pairs]
        return list(self._pairs)


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range when CUDA is available (a no-op otherwise)."""
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()
