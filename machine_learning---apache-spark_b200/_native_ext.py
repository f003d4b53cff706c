"""Signatures of the optional kernel groups (attention, comm, conv, lstm, mlp, ...).

Kept apart from ``_native.py`` so a library built before a group existed still loads; ``declare`` skips symbols
the loaded .so does not export.
"""
from __future__ import annotations

import ctypes as C

vp, i, ll, f, sz, u = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t, C.c_uint

SIGNATURES: dict[str, list] = {}
SIZE_T_RETURNS: dict[str, list] = {}


def register(name: str, argtypes: list, size_t_return: bool = False) -> None:
    (SIZE_T_RETURNS if size_t_return else SIGNATURES)[name] = argtypes


def declare(lib) -> None:
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.argtypes = args
            fn.restype = i
    for name, args in SIZE_T_RETURNS.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.argtypes = args
            fn.restype = sz
