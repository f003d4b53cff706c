"""Per-device scratch buffers for kernels that need global-memory workspace (two-stage reductions, split-K).

Buffers are cached by name and nothing is allocated in the steady state.  When a request outgrows a buffer a larger
one replaces it, but the superseded buffer is RETIRED, never freed: a CUDA graph captured earlier (the engine keeps one
per input shape, e.g. per LSTM sequence length) has the old address baked in, and replaying it must keep writing into
memory nobody else owns.  Growth is geometric, so the retired buffers of one name sum to less than twice its final
size.
"""
from __future__ import annotations

import torch

_cache: dict = {}
_retired: list = []   # superseded buffers that captured graphs may still reference
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
        size = max(int(nbytes), 256)
        if buf is not None:
            _retired.append(buf)
            size = max(size, buf.numel() * 3 // 2)
        buf = torch.empty(size, dtype=torch.uint8, device=device)
        _cache[key] = buf
    return buf


def get_zeroed(name: str, nbytes: int, device) -> torch.Tensor:
    """Like ``get``, but the buffer is zero-filled when it is (re)allocated - for kernels that keep counters in their
    scratch and leave them at zero after every call (e.g. the hot-row list of the embedding backward)."""
    device = torch.device(device)
    key = (name, device.type, device.index)
    before = _cache.get(key)
    buf = get(name, nbytes, device)
    if buf is not before:
        buf.zero_()
    return buf


def retired_bytes() -> int:
    return sum(b.numel() for b in _retired)


def clear() -> None:
    """Drop every buffer, retired ones included.  Only safe once no captured graph that used them will be replayed."""
    _cache.clear()
    _retired.clear()
