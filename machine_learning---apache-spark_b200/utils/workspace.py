"""Per-device scratch buffers for kernels that need global-memory workspace (two-stage reductions, split-K).

Buffers are cached by name and only ever grow, so their addresses are stable across steps — a requirement for
CUDA-graph replay — and nothing is allocated in the steady state.
"""
from __future__ import annotations

import torch

_cache: dict = {}


def get(name: str, nbytes: int, device) -> torch.Tensor:
    device = torch.device(device)
    key = (name, device.type, device.index)
    buf = _cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _cache[key] = buf
    return buf


def clear() -> None:
    _cache.clear()
