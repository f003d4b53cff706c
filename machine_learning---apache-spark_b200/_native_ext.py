"""Signatures of the kernel groups declared next to their Python ops (attention, comm, conv, lstm, mlp, ...).

``register`` may run before or after the library is loaded; ``declare`` skips symbols the loaded .so does not export.
"""
from __future__ import annotations

import ctypes as C

vp, i, ll, f, sz, u = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t, C.c_uint

SIGNATURES: dict[str, list] = {}
SIZE_T_RETURNS: dict[str, list] = {}


def _apply(lib, name, args, size_t_return):
    fn = getattr(lib, name, None)
    if fn is not None:
        fn.argtypes = args
        fn.restype = sz if size_t_return else i


def register(name: str, argtypes: list, size_t_return: bool = False) -> None:
    (SIZE_T_RETURNS if size_t_return else SIGNATURES)[name] = argtypes
    from . import _native

    if _native._lib is not None:
        _apply(_native._lib, name, argtypes, size_t_return)


def declare(lib) -> None:
    for name, args in SIGNATURES.items():
        _apply(lib, name, args, False)
    for name, args in SIZE_T_RETURNS.items():
        _apply(lib, name, args, True)
