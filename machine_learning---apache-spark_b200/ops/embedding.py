"""Embedding gather (+ sinusoidal positional encoding + dropout) and its backward (csrc/kernels/embedding.cu).
Reference: transformer.py:27-62; distributed_lstm.py:115,128.

Backward is deterministic: the batch's (token, position) keys are sorted once - at forward time, on an auxiliary stream,
because only the token ids are needed - and each row of the table gradient is then written exactly once by the warp
that owns it (sum over its tokens in position order, zero for rows the batch does not touch).  No clearing of the
table gradient, no atomics, the same bits on every run."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native, _native_ext
from ..utils import workspace
from ..utils.rng import DeviceCounters
from .gemm import working_copy

_vp, _i, _ll, _f, _u = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_uint
_native_ext.register("b200ml_embedding_index", [_vp, _i, _i, _vp, _vp, _vp])
_native_ext.register("b200ml_embedding_index_bytes", [_i, _i], size_t_return=True)
_native_ext.register("b200ml_embedding_bwd_sorted", [_vp, _vp, _i, _vp, _i, _i, _i, _ll, _f, _vp, _u, _i, _vp,
                                                     C.c_size_t, _vp])
_native_ext.register("b200ml_embedding_bwd_sorted_scratch_bytes", [_i, _i], size_t_return=True)
INDEX_MAX_TOKENS = 16384   # csrc/kernels/embedding.cu kIndexMaxTokens

_pe_cache: dict = {}
_aux_streams: dict = {}


def _aux_stream(device) -> torch.cuda.Stream:
    key = (device.type, device.index)
    st = _aux_streams.get(key)
    if st is None:
        st = _aux_streams[key] = torch.cuda.Stream(device=device)
    return st


def build_token_index(tok: torch.Tensor, V: int, touched=None):
    """(index, event): the backward index of one batch (sorted ``token << 32 | position`` keys, segment starts, hot
    segments, touched-row flags - csrc/kernels/embedding.cu IndexLayout), built by ONE CTA on the auxiliary stream;
    ``event`` marks its completion.  Only called for batches of at most INDEX_MAX_TOKENS tokens."""
    n = tok.numel()
    lib = _native.load()
    cur = torch.cuda.current_stream()
    aux = _aux_stream(tok.device)
    # the buffer is allocated on the CURRENT stream (it is read there, in backward, after waiting for `event`) and only
    # written on the auxiliary one; `tok` is saved for backward, so both outlive the auxiliary kernel
    index = torch.empty(int(lib.b200ml_embedding_index_bytes(n, V)) // 8 + 1, dtype=torch.int64, device=tok.device)
    aux.wait_stream(cur)
    with torch.cuda.stream(aux):
        # `touched` (optional): the table's peer-mapped touched-row flags, read by the other ranks' all-reduce kernels
        _native.check(lib.b200ml_embedding_index(tok.data_ptr(), n, V, index.data_ptr(), _native.ptr(touched),
                                                 _native.stream_ptr()), "embedding_index")
        ev = torch.cuda.Event()
        ev.record(aux)
    return index, ev


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
            keys = tok
            ctx.keys_event = None
            touched = getattr(weight, "_b200_touched", None)   # set by the DP engine for row-sparse all-reduce
            if weight.requires_grad and touched is not None and (
                    n > INDEX_MAX_TOKENS or getattr(weight, "_b200_grad_accumulate", False)):
                touched.fill_(1)      # atomicAdd path / accumulated gradients: every row may be non-zero
                touched = None
            if weight.requires_grad and n <= INDEX_MAX_TOKENS:
                keys, ctx.keys_event = build_token_index(tok, V, touched)
            ctx.save_for_backward(tok, keys)
            ctx.mask = None
            return out.reshape(*lead, d)
        out = weight.float()[tok]
        if pe is not None:
            out = (out.reshape(-1, S, d) + pe[:S].float()).reshape(n, d)
        mask = None
        if p > 0:
            mask = (torch.rand_like(out) >= p).float() / (1.0 - p)
            out = out * mask
        ctx.save_for_backward(tok, tok)
        ctx.mask = mask
        return out.to(out_dtype).reshape(*lead, d)

    @staticmethod
    def backward(ctx, dout):
        tok, keys = ctx.saved_tensors
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
            acc = 0
            if slot is not None:
                dt = slot.view(V, d)
                acc = int(bool(getattr(weight, "_b200_grad_accumulate", False)))
            else:
                dt = torch.empty(V, d, dtype=torch.float32, device=weight.device)
            rng = DeviceCounters.get(weight.device).rng.data_ptr() if p > 0 else None
            lib = _native.load()
            pad = -1 if padding_idx is None else padding_idx
            if ctx.keys_event is not None:
                # sorted path: every row is written exactly once (zeros where the batch has no token): nothing is
                # cleared beforehand, nothing is added atomically
                torch.cuda.current_stream().wait_event(ctx.keys_event)
                nb = int(lib.b200ml_embedding_bwd_sorted_scratch_bytes(n, d))
                # one scratch per table: the backward passes of two tables (encoder / decoder) may run concurrently
                ws = workspace.get_zeroed(f"emb_bwd@{id(weight)}", nb, weight.device)
                _native.check(lib.b200ml_embedding_bwd_sorted(
                    keys.data_ptr(), g.data_ptr(), int(g.dtype == torch.float32), dt.data_ptr(), n, d, V, pad, p, rng,
                    site, acc, ws.data_ptr(), ws.numel(), _native.stream_ptr()), "embedding_bwd_sorted")
            else:
                # more tokens than the single-CTA sort takes: clear + fp32 atomicAdd scatter (sum order not fixed)
                if not acc:
                    dt.zero_()
                _native.check(lib.b200ml_embedding_bwd(
                    tok.data_ptr(), g.data_ptr(), int(g.dtype == torch.float32), dt.data_ptr(), n, d, V, pad, p, rng,
                    site, _native.stream_ptr()), "embedding_bwd")
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
        if slot is not None:
            dt = slot.view(V, d)
            if not getattr(weight, "_b200_grad_accumulate", False):
                dt.zero_()
        else:
            dt = torch.zeros(V, d, dtype=torch.float32, device=weight.device)
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
