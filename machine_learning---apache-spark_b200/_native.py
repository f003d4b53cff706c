"""ctypes binding of libb200ml.so (C ABI in include/b200ml.h).

The library is built in-tree by ``build.py`` (nvcc, sm_100a).  On a machine with a CUDA device the library MUST
load — ops never fall back silently to eager PyTorch there; on a CPU-only machine ``available()`` is False and the
ops use their PyTorch restatements (which double as the numerics oracles in tests).
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200ml.so")

_lib = None
_lock = threading.Lock()
_launches = 0  # kernels launched through this binding (bench.py reports it as gpu_launches)


class NativeError(RuntimeError):
    pass


class GemmArgs(C.Structure):
    _fields_ = [
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("A", C.c_void_p), ("lda", C.c_longlong), ("a_mn_major", C.c_int),
        ("B", C.c_void_p), ("ldb", C.c_longlong), ("b_mn_major", C.c_int),
        ("C", C.c_void_p), ("ldc", C.c_longlong), ("c_fp32", C.c_int),
        ("bias", C.c_void_p),
        ("act", C.c_int),
        ("aux", C.c_void_p), ("ldaux", C.c_longlong), ("aux_mode", C.c_int), ("aux_scale", C.c_float),
        ("alpha", C.c_float),
        ("alpha_dev", C.c_void_p),
        ("accumulate", C.c_int),
        ("dropout_p", C.c_float), ("rng_state", C.c_void_p), ("rng_site", C.c_uint),
        ("tile_n", C.c_int),
        ("force_simt", C.c_int),
        ("split_k", C.c_int),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("epilogue_direct", C.c_int),
        ("max_ctas", C.c_int),
        ("row_stats", C.c_void_p), ("row_stats_ld", C.c_int),
    ]


def _declare(lib) -> None:
    vp, i, ll, f, sz, u = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t, C.c_uint
    lib.b200ml_last_error.restype = C.c_char_p
    lib.b200ml_version.restype = i
    sigs = {
        "b200ml_device_info": [vp, vp, vp, vp],
        "b200ml_gemm_bf16": [C.POINTER(GemmArgs), vp],
        "b200ml_layernorm_fwd": [vp, vp, vp, vp, vp, vp, vp, vp, i, i, f, f, vp, u, vp],
        "b200ml_layernorm_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, i, i, f, vp, u, i, vp],
        "b200ml_pe_table": [vp, i, i, vp],
        "b200ml_embedding_fwd": [vp, vp, vp, vp, i, i, i, i, i, f, vp, u, vp],
        "b200ml_embedding_bwd": [vp, vp, i, vp, i, i, i, ll, f, vp, u, vp],
        "b200ml_xent_count": [vp, i, ll, vp, vp],
        "b200ml_xent_fwd_bwd": [vp, i, ll, vp, vp, vp, f, i, i, ll, i, vp],
        "b200ml_loss_reduce": [vp, i, vp, f, vp, vp],
        "b200ml_optim_step": [vp, vp, i, vp, vp, vp, ll, i, f, f, f, f, f, vp, f, vp],
        "b200ml_cast_f32_bf16": [vp, vp, ll, vp],
        "b200ml_colsum_bf16": [vp, ll, i, i, vp, i, vp, sz, vp],
        "b200ml_counters_advance": [vp, vp, vp],
        "b200ml_argmax_accuracy": [vp, i, ll, vp, i, i, vp, vp, vp],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = args
        fn.restype = i
    for name, args in {"b200ml_layernorm_bwd_scratch_bytes": [i, i], "b200ml_colsum_scratch_bytes": [i, i]}.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = sz
    # optional groups (declared by their own modules once present)
    from . import _native_ext

    _native_ext.declare(lib)


def load():
    """Load (once) and return the ctypes library handle; raises NativeError when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise NativeError(
                    f"{LIB_PATH} is missing - run `python build.py` (or __graft_entry__.build()) first; "
                    "b200ml never falls back to eager PyTorch on a CUDA device")
            lib = C.CDLL(LIB_PATH)
            _declare(lib)
            _lib = lib
    return _lib


def available() -> bool:
    """True when kernels can run: CUDA device present (the library is then required to load)."""
    if not torch.cuda.is_available():
        return False
    load()
    return True


def use_native(*tensors) -> bool:
    """Dispatch rule: CUDA tensors -> native kernels (mandatory); CPU tensors -> PyTorch restatement."""
    for t in tensors:
        if t is not None and isinstance(t, torch.Tensor):
            if t.is_cuda:
                load()
                return True
            return False
    return False


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def check(rc: int, what: str = "") -> None:
    global _launches
    if rc != 0:
        msg = load().b200ml_last_error()
        raise NativeError(f"{what}: {msg.decode() if msg else 'unknown error'}")
    _launches += 1


def launches() -> int:
    return _launches


def ptr(t) -> int | None:
    return None if t is None else t.data_ptr()
