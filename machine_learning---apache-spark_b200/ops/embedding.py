"""Embedding gather (+ sinusoidal positional encoding + dropout) and its dense scatter-add backward
(csrc/kernels/embedding.cu).  Reference: transformer.py:27-62; distributed_lstm.py:115,128."""
from __future__ import annotations

import torch

from .. import _native
from ..utils.rng import DeviceCounters
from .gemm import working_copy

_pe_cache: dict = {}


def positional_encoding_table(S: int, d: int, device) -> torch.Tensor:
    """[S, d] fp32 table, PE[pos,2i]=sin(pos/10000^(2i/d)), PE[pos,2i+1]=cos(..) — built once per (S, d, device)
    instead of on the CPU every forward as transformer.py:33-42 does."""
    device = torch.device(device)
    key = (S, d, device.type, device.index)
    pe = _pe_cache.get(key)
    if pe is None:
        if device.type == "cuda":
            pe = torch.empty(S, d, dtype=torch.float32, device=device)
            _native.check(_native.load().b200ml_pe_table(pe.data_ptr(), S, d, _native.stream_ptr()), "pe_table")
        else:
            even_i = torch.arange(0, d, 2).float()
            denom = torch.pow(10000, even_i / d)
            pos = torch.arange(S).reshape(S, 1)
            pe = torch.stack([torch.sin(pos / denom), torch.cos(pos / denom)], dim=2).flatten(1, 2)
        _pe_cache[key] = pe
    return pe


class _Embedding(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens, weight, pe, p, site, padding_idx, out_dtype):
        V, d = weight.shape
        lead = tokens.shape
        S = lead[-1]
        tok = tokens.reshape(-1).contiguous()
        n = tok.numel()
        ctx.meta = (V, d, p, site, padding_idx, lead)
        ctx.weight = weight
        if _native.use_native(weight):
            w16 = working_copy(weight)
            out = torch.empty(n, d, dtype=out_dtype, device=weight.device)
            rng = DeviceCounters.get(weight.device).rng.data_ptr() if p > 0 else None
            _native.check(_native.load().b200ml_embedding_fwd(
                tok.data_ptr(), w16.data_ptr(), _native.ptr(pe), out.data_ptr(), int(out_dtype == torch.float32), n, S,
                d, V, p, rng, site, _native.stream_ptr()), "embedding_fwd")
            ctx.save_for_backward(tok)
            ctx.mask = None
            return out.reshape(*lead, d)
        out = weight.float()[tok]
        if pe is not None:
            out = (out.reshape(-1, S, d) + pe[:S].float()).reshape(n, d)
        mask = None
        if p > 0:
            mask = (torch.rand_like(out) >= p).float() / (1.0 - p)
            out = out * mask
        ctx.save_for_backward(tok)
        ctx.mask = mask
        return out.to(out_dtype).reshape(*lead, d)

    @staticmethod
    def backward(ctx, dout):
        (tok,) = ctx.saved_tensors
        V, d, p, site, padding_idx, lead = ctx.meta
        weight = ctx.weight
        n = tok.numel()
        if not weight.requires_grad:
            return (None,) * 7
        slot = getattr(weight, "_b200_grad", None)
        if _native.use_native(weight):
            g = dout.reshape(n, d).contiguous()
            if g.dtype not in (torch.float32, torch.bfloat16):
                g = g.float()
            if slot is not None:
                dt = slot.view(V, d)  # zeroed by the engine at the start of the step
            else:
                dt = torch.zeros(V, d, dtype=torch.float32, device=weight.device)
            rng = DeviceCounters.get(weight.device).rng.data_ptr() if p > 0 else None
            _native.check(_native.load().b200ml_embedding_bwd(
                tok.data_ptr(), g.data_ptr(), int(g.dtype == torch.float32), dt.data_ptr(), n, d, V,
                -1 if padding_idx is None else padding_idx, p, rng, site, _native.stream_ptr()), "embedding_bwd")
            if slot is not None:
                ready = getattr(weight, "_b200_ready", None)
                if ready is not None:
                    ready()
                return (None,) * 7
            return None, dt, None, None, None, None, None
        g = dout.reshape(n, d).float()
        if ctx.mask is not None:
            g = g * ctx.mask
        if padding_idx is not None:
            g = g * (tok != padding_idx).unsqueeze(1)
        dt = slot.view(V, d) if slot is not None else torch.zeros(V, d, dtype=torch.float32, device=weight.device)
        dt.index_add_(0, tok, g)
        if slot is not None:
            ready = getattr(weight, "_b200_ready", None)
            if ready is not None:
                ready()
            return (None,) * 7
        return None, dt.to(weight.dtype), None, None, None, None, None


def embedding(tokens, weight, pe=None, dropout_p: float = 0.0, site: int = 0, padding_idx=None,
              out_dtype=torch.bfloat16, training: bool = True):
    """dropout(weight[tokens] + pe[pos]) with pos = index along the last token dimension."""
    return _Embedding.apply(tokens, weight, pe, dropout_p if training else 0.0, site, padding_idx, out_dtype)
