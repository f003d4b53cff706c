"""Device-resident RNG + step counters.

Dropout masks are a pure function of (seed, step, site, element index) — csrc/common/rng.cuh — so backward
regenerates them and a CUDA-graph replay sees fresh masks: ``step`` lives in device memory and is bumped by a
one-thread kernel (``advance``) that is part of the captured step.
"""
from __future__ import annotations

import itertools

import torch

from .. import _native

_site_counter = itertools.count(1)


def new_site() -> int:
    """A process-unique id for one dropout call site (same value on every rank: allocation order is deterministic)."""
    return next(_site_counter)


class DeviceCounters:
    """{uint64 seed, uint64 step} for dropout + an int32 optimizer step, all on the device."""

    _per_device: dict = {}

    def __init__(self, device, seed: int = 0):
        self.device = torch.device(device)
        self.rng = torch.tensor([seed, 0], dtype=torch.int64, device=self.device)
        self.opt_step = torch.zeros(1, dtype=torch.int32, device=self.device)

    @classmethod
    def get(cls, device) -> "DeviceCounters":
        device = torch.device(device)
        key = (device.type, device.index if device.index is not None else (torch.cuda.current_device() if device.type == "cuda" else 0))
        if key not in cls._per_device:
            cls._per_device[key] = cls(device)
        return cls._per_device[key]

    def seed(self, seed: int) -> None:
        self.rng[0] = seed
        self.rng[1] = 0

    def advance(self, rng: bool = True, opt: bool = False, opt_tensor=None) -> None:
        """One one-thread kernel bumps the dropout step and (optionally) an optimizer step counter — the caller's own
        ``opt_tensor`` (each engine/optimizer counts its own steps) or the shared ``opt_step``."""
        step = opt_tensor if opt_tensor is not None else (self.opt_step if opt else None)
        if self.device.type == "cuda":
            lib = _native.load()
            _native.check(lib.b200ml_counters_advance(self.rng.data_ptr() if rng else None,
                                                      None if step is None else step.data_ptr(),
                                                      _native.stream_ptr()), "counters_advance")
        else:
            if rng:
                self.rng[1] += 1
            if step is not None:
                step += 1


def manual_seed(seed: int, device=None) -> None:
    """Seed both torch and the b200ml dropout stream (the reference never seeds; tests need to)."""
    torch.manual_seed(seed)
    if device is None:
        device = "cuda" if torch.cuda.is_available() else "cpu"
    DeviceCounters.get(device).seed(seed)
