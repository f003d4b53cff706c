"""Gradient bucket manager — ctypes face of csrc/runtime/reducer.cpp (the role of torch's c10d Reducer behind
``DistributedDataParallel``: distributed_cnn.py:156, distributed_lstm.py:160, distributed_multilayer_perceptron.py:106).

Buckets are contiguous ranges of the flat gradient buffer; ``mark_ready(i)`` returns the bucket index when parameter
``i`` was the last one its bucket waited for.  ``PyReducer`` is the same contract in Python, used only when the native
library has not been built (a CPU-only checkout without nvcc).
"""
from __future__ import annotations

import ctypes as C

from .. import _native, _native_ext

_vp, _i, _ll = C.c_void_p, C.c_int, C.c_longlong
_native_ext.register("b200ml_reducer_num_buckets", [_vp])
_native_ext.register("b200ml_reducer_bucket", [_vp, _i, _vp, _vp, _vp, _vp, _vp])
_native_ext.register("b200ml_reducer_bucket_of", [_vp, _i])
_native_ext.register("b200ml_reducer_mark_ready", [_vp, _i])
_native_ext.register("b200ml_reducer_flush", [_vp, _vp, _i])


class NativeReducer:
    def __init__(self, offsets: list[int], group_ends: list[int], cap_elems: int):
        lib = _native.load()
        lib.b200ml_reducer_create.restype = _vp
        lib.b200ml_reducer_create.argtypes = [_i, _vp, _i, _vp, _ll]
        lib.b200ml_reducer_destroy.restype = None
        lib.b200ml_reducer_destroy.argtypes = [_vp]
        lib.b200ml_reducer_begin_step.restype = None
        lib.b200ml_reducer_begin_step.argtypes = [_vp]
        self._lib = lib
        n = len(offsets) - 1
        off = (_ll * (n + 1))(*offsets)
        ge = (_i * len(group_ends))(*group_ends)
        self._h = lib.b200ml_reducer_create(n, off, len(group_ends), ge, int(cap_elems))
        if not self._h:
            msg = lib.b200ml_last_error()
            raise _native.NativeError(f"reducer_create: {msg.decode() if msg else 'failed'}")
        self.buckets = []   # (start, numel, [param indices], group)
        for b in range(lib.b200ml_reducer_num_buckets(self._h)):
            s, m, f, c, g = _ll(), _ll(), _i(), _i(), _i()
            lib.b200ml_reducer_bucket(self._h, b, C.byref(s), C.byref(m), C.byref(f), C.byref(c), C.byref(g))
            self.buckets.append((s.value, m.value, list(range(f.value, f.value + c.value)), g.value))
        self._flush_buf = (_i * max(len(self.buckets), 1))()

    def bucket_of(self, param: int) -> int:
        return self._lib.b200ml_reducer_bucket_of(self._h, param)

    def begin_step(self) -> None:
        self._lib.b200ml_reducer_begin_step(self._h)

    def mark_ready(self, param: int) -> int:
        return self._lib.b200ml_reducer_mark_ready(self._h, param)

    def flush(self) -> list[int]:
        n = self._lib.b200ml_reducer_flush(self._h, self._flush_buf, len(self._flush_buf))
        return [self._flush_buf[k] for k in range(min(n, len(self._flush_buf)))]

    def close(self) -> None:
        if self._h:
            self._lib.b200ml_reducer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PyReducer:
    def __init__(self, offsets: list[int], group_ends: list[int], cap_elems: int):
        self.buckets, self._of = [], {}
        lo = 0
        for g, hi in enumerate(group_ends):
            start, first, big_end = (offsets[lo] if lo < hi else 0), lo, -1

            def is_big(i):
                return (offsets[i + 1] - offsets[i]) * 4 >= cap_elems

            for i in range(lo, hi):          # same rules as reducer.cpp
                if is_big(i):
                    big_end = offsets[i + 1]
                end = offsets[i + 1]
                close = i + 1 == hi
                if not close:
                    if is_big(i + 1):
                        close = True
                    elif big_end >= 0:
                        close = offsets[i + 2] - big_end > cap_elems // 2
                    else:
                        close = end - start >= cap_elems
                if close:
                    for p in range(first, i + 1):
                        self._of[p] = len(self.buckets)
                    self.buckets.append((start, end - start, list(range(first, i + 1)), g))
                    start, first, big_end = end, i + 1, -1
            lo = hi
        self._pending = [0] * len(self.buckets)
        self._reported: set = set()

    def bucket_of(self, param: int) -> int:
        return self._of.get(param, -1)

    def begin_step(self) -> None:
        self._pending = [len(b[2]) for b in self.buckets]
        self._reported = set()

    def mark_ready(self, param: int) -> int:
        if param in self._reported:
            return -1
        self._reported.add(param)
        b = self._of[param]
        self._pending[b] -= 1
        return b if self._pending[b] == 0 else -1

    def flush(self) -> list[int]:
        out = [b for b, p in enumerate(self._pending) if p > 0]
        for b in out:
            self._pending[b] = 0
        return out

    def close(self) -> None:
        pass


def make_reducer(offsets, group_ends, cap_elems):
    try:
        return NativeReducer(offsets, group_ends, cap_elems)
    except (_native.NativeError, OSError):
        return PyReducer(offsets, group_ends, cap_elems)
