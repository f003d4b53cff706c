"""Autograd wrapper of the fused conv3x3 + bias + ReLU (+ max-pool) kernels (csrc/kernels/conv.cu)."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native, _native_ext
from ..utils import workspace
from .gemm import deliver_grad

vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
_native_ext.register("b200ml_conv_set_algo", [i])
_native_ext.register("b200ml_conv3x3_relu_fwd", [vp, vp, vp, vp, vp, i, i, i, i, i, i, vp])
_native_ext.register("b200ml_conv3x3_bwd_scratch_bytes", [i, i, i, i, i], size_t_return=True)
_native_ext.register("b200ml_conv3x3_relu_bwd", [vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, i, i, i, i, i, i, i, vp])


class _ConvReluPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, pool):
        lib = _native.load()
        x = x.float().contiguous()
        B, Cin, H, W = x.shape
        Cout = weight.shape[0]
        w = weight.detach().float().contiguous()
        b = None if bias is None else bias.detach().float().contiguous()
        Ho, Wo = (H // 2, W // 2) if pool else (H, W)
        y = torch.empty(B, Cout, Ho, Wo, dtype=torch.float32, device=x.device)
        idx = torch.empty(B, Cout, Ho, Wo, dtype=torch.uint8, device=x.device) if pool else None
        _native.check(lib.b200ml_conv3x3_relu_fwd(x.data_ptr(), w.data_ptr(), _native.ptr(b), y.data_ptr(),
                                                  _native.ptr(idx), B, Cin, Cout, H, W, int(pool),
                                                  _native.stream_ptr()), "conv3x3_relu_fwd")
        ctx.params = (weight, bias)
        ctx.pool = pool
        ctx.save_for_backward(x, w, y, idx if pool else y.new_empty(0, dtype=torch.uint8))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y, idx = ctx.saved_tensors
        weight, bias = ctx.params
        lib = _native.load()
        B, Cin, H, W = x.shape
        Cout = w.shape[0]
        dy = dy.float().contiguous()
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        nbytes = lib.b200ml_conv3x3_bwd_scratch_bytes(B, Cin, Cout, H, W)
        ws = workspace.get("conv_bwd", nbytes, x.device)
        slot_w, slot_b = getattr(weight, "_b200_grad", None), None if bias is None else getattr(bias, "_b200_grad", None)
        direct = slot_w is not None and (bias is None or slot_b is not None)
        if direct:
            dw, db = slot_w, (slot_b if bias is not None else torch.empty(Cout, dtype=torch.float32, device=x.device))
            acc = int(bool(getattr(weight, "_b200_grad_accumulate", False)))
        else:
            dw = torch.empty(Cout * Cin * 9, dtype=torch.float32, device=x.device)
            db = torch.empty(Cout, dtype=torch.float32, device=x.device)
            acc = 0
        _native.check(lib.b200ml_conv3x3_relu_bwd(dy.data_ptr(), y.data_ptr(), idx.data_ptr() if ctx.pool else None,
                                                  x.data_ptr(), w.data_ptr(), _native.ptr(dx), dw.data_ptr(),
                                                  db.data_ptr(), ws.data_ptr(), ws.numel(), B, Cin, Cout, H, W,
                                                  int(ctx.pool), acc, _native.stream_ptr()), "conv3x3_relu_bwd")
        if direct:
            for prm in (weight, bias):
                if prm is not None:
                    ready = getattr(prm, "_b200_ready", None)
                    if ready is not None:
                        ready()
            return dx, None, None, None
        return dx, dw.view_as(weight), (db if bias is not None else None), None


def conv3x3_relu(x, weight, bias, pool: bool):
    assert weight.shape[2:] == (3, 3), "only 3x3 kernels (stride 1, padding 1) are implemented"
    return _ConvReluPool.apply(x, weight, bias, pool)


def set_algo(algo: str) -> None:
    """'auto' (tcgen05 implicit GEMM for forward / data gradient once 9*C >= 64), 'direct' (fp32 CUDA-core kernels) or
    'igemm' (always the implicit GEMM, bf16 operands).  Also the B200ML_CONV_ALGO environment variable (0/1/2)."""
    _native.load().b200ml_conv_set_algo({"auto": 0, "direct": 1, "igemm": 2}[algo])
