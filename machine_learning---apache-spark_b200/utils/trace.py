"""Tracing / profiling helpers (the reference only has wall-clock prints).

* ``nvtx_range(name)``      NVTX push/pop (shows up in ncu/nsys timelines), no-op on CPU
* ``PhaseTimer``            CUDA-event timers per named phase (forward / backward / comm / optimizer), read lazily so
                            the timed region is never synchronised
* ``kernel_launches()``     how many b200ml kernels this process has launched (bench.py's ``gpu_launches``)
"""
from __future__ import annotations

import contextlib
import time
from collections import defaultdict

import torch

from .. import _native


@contextlib.contextmanager
def nvtx_range(name: str):
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class PhaseTimer:
    def __init__(self, device="cuda"):
        self.cuda = torch.device(device).type == "cuda" and torch.cuda.is_available()
        self._pairs = defaultdict(list)
        self._cpu = defaultdict(float)

    @contextlib.contextmanager
    def phase(self, name: str):
        if self.cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            with nvtx_range(name):
                yield
            e1.record()
            self._pairs[name].append((e0, e1))
        else:
            t0 = time.perf_counter()
            yield
            self._cpu[name] += (time.perf_counter() - t0) * 1e3

    def summary(self) -> dict:
        """Milliseconds per phase (synchronises once, here)."""
        out = dict(self._cpu)
        if self.cuda:
            torch.cuda.synchronize()
            for name, pairs in self._pairs.items():
                out[name] = out.get(name, 0.0) + sum(a.elapsed_time(b) for a, b in pairs)
        return out

    def reset(self) -> None:
        self._pairs.clear()
        self._cpu.clear()


def kernel_launches() -> int:
    return _native.launches()
