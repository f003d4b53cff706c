"""Per-device scratch buffers for kernels that need global-memory workspace (two-stage reductions, split-K).

Buffers are cached by name and only ever grow, so their addresses are stable across steps — a requirement for
CUDA-graph replay — and nothing is allocated in the steady state.
"""
from __future__ import annotations

import torch

_cache: dict = {}
_lanes: dict = {}   # cuda stream pointer -> small lane id


def register_stream(stream, lane: int) -> None:
    """Give a long-lived side stream its own set of scratch buffers.  Streams that are never registered (the default
    stream, a CUDA-graph capture stream) share lane 0 — they never run concurrently with each other."""
    _lanes[stream.cuda_stream] = lane


def lane() -> int:
    return _lanes.get(torch.cuda.current_stream().cuda_stream, 0) if torch.cuda.is_available() else 0


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
