"""Multi-head attention on the tcgen05 kernels (csrc/kernels/attention.cu), consuming the per-head interleaved
qkv / kv Linear outputs in place.

Reference: transformer.py:12-25 (scaled_dot_product), :65-83 (MultiHeadAttention: qkv_layer output reshaped
[B,S,H,3·hd] → q,k,v = chunk(3)), :169-191 (MultiHeadCrossAttention: kv_layer on the encoder output [B,S,H,2·hd],
q_layer on the decoder stream).  ``mask_mode``:
  "reference" — the boolean look-ahead mask is ADDED to the scores after a transpose (True → +1.0), i.e. +1.0 where
                key < query; this is what transformer.py:17-20 computes as written.  The encoder's
                padding mask adds a per-query constant and cancels in the softmax, so "none" reproduces it.
  "causal"    — -inf where key > query (the intended behaviour).
  "none"      — no bias.
"""
from __future__ import annotations

import ctypes as C
import math
import warnings

import torch

from .. import _native, _native_ext

MASK_MODES = {"none": 0, None: 0, "reference": 1, "causal": 2, 0: 0, 1: 1, 2: 2}


class AttnOperand(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("ld", C.c_longlong), ("col0", C.c_int), ("head_stride", C.c_int)]


_P = C.POINTER(AttnOperand)
_native_ext.register("b200ml_attn_fwd", [_P, _P, _P, _P, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                         C.c_int, C.c_void_p])
_native_ext.register("b200ml_attn_bwd", [_P, _P, _P, _P, _P, C.c_void_p, _P, _P, _P, C.c_int, C.c_int, C.c_int,
                                         C.c_int, C.c_float, C.c_int, C.c_void_p])


def _op(t, col0, head_stride):
    return AttnOperand(t.data_ptr(), t.stride(-2), col0, head_stride)


def _layout(kind, hd):
    # (q col0, q stride), (k col0, k stride), (v col0, v stride) inside their source matrices
    if kind == "self":
        return (0, 3 * hd), (hd, 3 * hd), (2 * hd, 3 * hd)
    return (0, hd), (0, 2 * hd), (hd, 2 * hd)


def _split(kind, qsrc, kvsrc, H, hd):
    """Views q,k,v [B,H,S,hd] from the interleaved sources (used by the PyTorch restatement)."""
    B, Sq = qsrc.shape[0], qsrc.shape[1]
    if kind == "self":
        t = qsrc.reshape(B, Sq, H, 3 * hd).permute(0, 2, 1, 3)
        return t[..., :hd], t[..., hd:2 * hd], t[..., 2 * hd:]
    Skv = kvsrc.shape[1]
    q = qsrc.reshape(B, Sq, H, hd).permute(0, 2, 1, 3)
    kv = kvsrc.reshape(B, Skv, H, 2 * hd).permute(0, 2, 1, 3)
    return q, kv[..., :hd], kv[..., hd:]


def _bias(mask_mode, Sq, Skv, device):
    if mask_mode == 0:
        return None
    qi = torch.arange(Sq, device=device)[:, None]
    ki = torch.arange(Skv, device=device)[None, :]
    if mask_mode == 1:
        return (ki < qi).float()
    return torch.zeros(Sq, Skv, device=device).masked_fill(ki > qi, float("-inf"))


def attention_reference(kind, qsrc, kvsrc, H, mask_mode):
    """fp32 PyTorch restatement; differentiable (used on CPU and as the test oracle)."""
    d = qsrc.shape[-1] // (3 if kind == "self" else 1)
    hd = d // H
    q, k, v = _split(kind, qsrc.float(), None if kvsrc is None else kvsrc.float(), H, hd)
    s = q @ k.transpose(-1, -2) / math.sqrt(hd)
    b = _bias(MASK_MODES[mask_mode], q.shape[2], k.shape[2], qsrc.device)
    if b is not None:
        s = s + b
    o = torch.softmax(s, -1) @ v
    B, _, Sq, _ = o.shape
    return o.permute(0, 2, 1, 3).reshape(B, Sq, H * hd)


_warned = False


class _Attention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qsrc, kvsrc, kind, H, mask_mode):
        B, Sq = qsrc.shape[0], qsrc.shape[1]
        d = qsrc.shape[-1] // (3 if kind == "self" else 1)
        hd = d // H
        Skv = Sq if kind == "self" else kvsrc.shape[1]
        ctx.meta = (kind, H, mask_mode, B, Sq, Skv, hd)
        lib = _native.load()
        qsrc = qsrc.contiguous()
        ksrc = qsrc if kind == "self" else kvsrc.contiguous()
        (qc, qs), (kc, ks), (vc, vs) = _layout(kind, hd)
        out = torch.empty(B, Sq, H * hd, dtype=torch.bfloat16, device=qsrc.device)
        lse = torch.empty(B, H, Sq, dtype=torch.float32, device=qsrc.device)
        _native.check(lib.b200ml_attn_fwd(C.byref(_op(qsrc, qc, qs)), C.byref(_op(ksrc, kc, ks)),
                                          C.byref(_op(ksrc, vc, vs)), C.byref(_op(out, 0, hd)), lse.data_ptr(), B, H,
                                          Sq, Skv, 1.0 / math.sqrt(hd), mask_mode, _native.stream_ptr()), "attn_fwd")
        ctx.save_for_backward(qsrc, ksrc, out, lse)
        return out

    @staticmethod
    def backward(ctx, dout):
        qsrc, ksrc, out, lse = ctx.saved_tensors
        kind, H, mask_mode, B, Sq, Skv, hd = ctx.meta
        lib = _native.load()
        dout = dout.to(torch.bfloat16).contiguous()
        (qc, qs), (kc, ks), (vc, vs) = _layout(kind, hd)
        dq_src = torch.empty_like(qsrc)
        dk_src = dq_src if kind == "self" else torch.empty_like(ksrc)
        _native.check(lib.b200ml_attn_bwd(
            C.byref(_op(qsrc, qc, qs)), C.byref(_op(ksrc, kc, ks)), C.byref(_op(ksrc, vc, vs)),
            C.byref(_op(out, 0, hd)), C.byref(_op(dout, 0, hd)), lse.data_ptr(), C.byref(_op(dq_src, qc, qs)),
            C.byref(_op(dk_src, kc, ks)), C.byref(_op(dk_src, vc, vs)), B, H, Sq, Skv, 1.0 / math.sqrt(hd),
            mask_mode, _native.stream_ptr()), "attn_bwd")
        return dq_src, (None if kind == "self" else dk_src), None, None, None


def attention(qsrc, kvsrc=None, *, num_heads: int, mask_mode="none"):
    """Self-attention (``kvsrc`` None: ``qsrc`` = qkv_layer output [B,S,3d]) or cross-attention (``qsrc`` = q_layer
    output [B,Sq,d], ``kvsrc`` = kv_layer output [B,Skv,2d]).  Returns [B,Sq,d]."""
    global _warned
    kind = "self" if kvsrc is None else "cross"
    mm = MASK_MODES[mask_mode]
    d = qsrc.shape[-1] // (3 if kind == "self" else 1)
    hd = d // num_heads
    Sq = qsrc.shape[1]
    Skv = Sq if kvsrc is None else kvsrc.shape[1]
    native_ok = hd == 64 and Sq <= 256 and Skv <= 256
    if _native.use_native(qsrc) and native_ok:
        qs = qsrc if qsrc.dtype == torch.bfloat16 else qsrc.to(torch.bfloat16)
        ks = None if kvsrc is None else (kvsrc if kvsrc.dtype == torch.bfloat16 else kvsrc.to(torch.bfloat16))
        return _Attention.apply(qs, ks, kind, num_heads, mm)
    if qsrc.is_cuda and not _warned:
        warnings.warn(f"b200ml attention: head_dim={hd}, S=({Sq},{Skv}) is outside the tcgen05 kernel's envelope "
                      "(head_dim 64, S <= 256); using the PyTorch restatement")
        _warned = True
    return attention_reference(kind, qsrc, kvsrc, num_heads, mm).to(qsrc.dtype)
