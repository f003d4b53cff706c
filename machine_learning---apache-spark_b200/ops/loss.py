"""Softmax cross-entropy (csrc/kernels/xent.cu): a generic ``cross_entropy`` for the small classifier heads and
``linear_cross_entropy`` — the translator's 512→V_de projection, loss and both of its gradients as one autograd node
that never materialises fp32 logits or a separate gradient tensor.

Reference: nn.CrossEntropyLoss() (distributed_cnn.py:141, distributed_lstm.py:142,
distributed_multilayer_perceptron.py:109); CrossEntropyLoss(ignore_index=0, reduction='none') + masked mean
(pytorch_machine_translator.py:125,182-188); the head Linear (transformer.py:272,283).
"""
from __future__ import annotations

import os

import torch

import ctypes as C

from .. import _native, _native_ext
from ..utils import workspace
from .gemm import colsum, deliver_grad, gemm, working_copy

_vp, _i, _ll, _f, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t
_native_ext.register("b200ml_xent_rows", [_vp, _ll, _vp, _vp, _vp, _f, _i, _i, _ll, _i, _vp, _i, _vp, _sz, _vp])
_native_ext.register("b200ml_xent_rows_scratch_bytes", [_i], size_t_return=True)
_native_ext.register("b200ml_xent_rows_lse", [_vp, _ll, _vp, _vp, _vp, _f, _i, _i, _ll, _i, _vp, _i, _vp, _sz, _vp, _vp])
_native_ext.register("b200ml_xent_lse", [_vp, _ll, _i, _i, _vp, _vp])
# Softmax statistics from the head GEMM's epilogue (one pass of the loss kernel instead of three): needs the CTA-pair
# tile (256 rows) and is switched off with B200ML_HEAD_STATS=0
HEAD_STATS = os.environ.get("B200ML_HEAD_STATS", "1") != "0"
XENT_ROWS_MAX_V = 20480  # csrc/kernels/xent.cu kBigMaxV

# Optional row-chunking of the head (B200ML_HEAD_CHUNK_MB > 0) keeps each logits chunk L2-resident between the GEMM, the
# loss kernel and the input-gradient GEMM.  First measurement on B200 at the translator shape: it LOST (2.12 ms vs
# 1.80 ms per step at 32 MB chunks) because the per-chunk dgrad had too few output tiles to fill 148 SMs; see
# BASELINE.md for the re-measurement with the CTA-pair GEMM.  Default: one chunk.
HEAD_CHUNK_BYTES = int(float(os.environ.get("B200ML_HEAD_CHUNK_MB", "0")) * (1 << 20))
# The head's weight gradient is a leaf of backward but, as a persistent GEMM on every SM, it keeps the kernels of the
# critical path (LayerNorm backward, the decoder's dgrads) waiting for an SM.  > 0: run it on at most this many CTAs.
HEAD_WGRAD_CTAS = int(os.environ.get("B200ML_HEAD_WGRAD_CTAS", "0"))


def _xent_native(logits, ld, targets, V, ignore_index, write_grad=True):
    """Runs the count + fused fwd/bwd + reduce kernels.  ``logits`` is overwritten with the gradient."""
    lib = _native.load()
    rows = targets.numel()
    dev = logits.device
    inv = torch.empty(2, dtype=torch.float32, device=dev)
    row_loss = torch.empty(rows, dtype=torch.float32, device=dev)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    s = _native.stream_ptr()
    _native.check(lib.b200ml_xent_count(targets.data_ptr(), rows, ignore_index, inv.data_ptr(), s), "xent_count")
    _native.check(lib.b200ml_xent_fwd_bwd(logits.data_ptr(), int(logits.dtype == torch.float32), ld,
                                          targets.data_ptr(), row_loss.data_ptr(), inv.data_ptr(), 0.0, rows, V,
                                          ignore_index, int(write_grad), s), "xent_fwd_bwd")
    _native.check(lib.b200ml_loss_reduce(row_loss.data_ptr(), rows, inv.data_ptr(), 0.0, loss.data_ptr(), s),
                  "loss_reduce")
    return loss[0], row_loss


class _CrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, ignore_index):
        V = logits.shape[-1]
        z = logits.reshape(-1, V)
        t = targets.reshape(-1).contiguous()
        ctx.shape, ctx.dtype = logits.shape, logits.dtype
        if _native.use_native(z):
            g = z.detach().clone() if z.dtype in (torch.float32, torch.bfloat16) else z.detach().float()
            g = g.contiguous()
            loss, _ = _xent_native(g, g.stride(0), t, V, ignore_index)
            ctx.save_for_backward(g)
            return loss
        zf = z.float()
        valid = t != ignore_index
        lse = torch.logsumexp(zf, -1)
        picked = zf.gather(1, t.clamp(0, V - 1).unsqueeze(1)).squeeze(1)
        cnt = valid.sum().clamp(min=1)
        loss = ((lse - picked) * valid).sum() / cnt
        g = (torch.softmax(zf, -1) - torch.nn.functional.one_hot(t.clamp(0, V - 1), V)) * valid.unsqueeze(1) / cnt
        ctx.save_for_backward(g)
        return loss

    @staticmethod
    def backward(ctx, gout):
        (g,) = ctx.saved_tensors
        return (g.float() * gout).to(ctx.dtype).reshape(ctx.shape), None, None


def cross_entropy(logits, targets, ignore_index: int = -100):
    """Mean softmax cross-entropy over the rows whose target != ignore_index."""
    return _CrossEntropy.apply(logits, targets, ignore_index)


class _LinearCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, weight, bias, targets, ignore_index):
        V, d = weight.shape
        h2 = h.reshape(-1, d)
        t = targets.reshape(-1).contiguous()
        rows = h2.shape[0]
        ctx.params = (weight, bias)
        ctx.h_shape, ctx.h_dtype = h.shape, h.dtype
        if _native.use_native(h2):
            h2 = h2.to(torch.bfloat16).contiguous()
            w16 = working_copy(weight)
            ld = (V + 7) // 8 * 8  # TMA needs 16-byte row pitch
            dev = h.device
            lib = _native.load()
            st = _native.stream_ptr()
            logits = torch.empty(rows, ld, dtype=torch.bfloat16, device=dev)
            inv = torch.empty(2, dtype=torch.float32, device=dev)
            row_loss = torch.empty(rows, dtype=torch.float32, device=dev)
            loss = torch.empty(1, dtype=torch.float32, device=dev)
            need_grad = h.requires_grad or weight.requires_grad
            _native.check(lib.b200ml_xent_count(t.data_ptr(), rows, ignore_index, inv.data_ptr(), st), "xent_count")
            # Row chunks sized to stay L2-resident (126 MB): the GEMM writes a chunk of logits, the loss kernel reads it
            # twice and overwrites it with the gradient, and the input-gradient GEMM consumes that gradient — all before
            # the chunk is evicted, so the [rows, V] matrix makes one trip to HBM instead of four.
            chunk_rows = max(128, (HEAD_CHUNK_BYTES // (ld * 2)) // 128 * 128) if HEAD_CHUNK_BYTES > 0 else rows
            # the input gradient is computed here only in chunked mode (while the gradient chunk is L2-resident);
            # otherwise backward computes it with the upstream gradient folded into the GEMM epilogue
            dh = (torch.empty(rows, d, dtype=torch.bfloat16, device=dev)
                  if (need_grad and h.requires_grad and HEAD_CHUNK_BYTES > 0) else None)
            b_ = None if bias is None else bias.detach()
            # single-chunk large-vocabulary path: one read of the logits, gradient in place, bias gradient for free
            fused_rows = V <= XENT_ROWS_MAX_V
            db_pre = None
            if fused_rows and need_grad and bias is not None and bias.requires_grad:
                db_pre = torch.empty(ld, dtype=torch.float32, device=dev)[:V]  # the kernel writes V rounded up to 8
            # single chunk, wide vocabulary, enough rows for the 256-row pair tile: the GEMM epilogue also emits each
            # row's (max, sum-exp) per half tile, a small kernel merges them into lse, and the loss kernel makes ONE pass
            use_stats = HEAD_STATS and fused_rows and chunk_rows >= rows and rows >= 256 and V >= 512
            for r0 in range(0, rows, chunk_rows):
                r1 = min(rows, r0 + chunk_rows)
                lg = logits[r0:r1]
                lse = None
                if use_stats:
                    parts = 2 * ((V + 255) // 256)
                    stats = torch.empty(rows, parts, 2, dtype=torch.float32, device=dev)
                    gemm(h2, w16, out=lg, bias=b_, n_logical=V, tile_n=512, split_k=1, row_stats=stats)
                    lse = torch.empty(rows, dtype=torch.float32, device=dev)
                    _native.check(lib.b200ml_xent_lse(stats.data_ptr(), parts, parts, rows, lse.data_ptr(), st), "xent_lse")
                else:
                    gemm(h2[r0:r1], w16, out=lg, bias=b_, n_logical=V)
                if fused_rows:
                    nb = int(lib.b200ml_xent_rows_scratch_bytes(V)) if db_pre is not None else 0
                    ws = workspace.get(f"xent_rows@{workspace.lane()}", nb, dev) if nb else None
                    if lse is not None:
                        _native.check(lib.b200ml_xent_rows_lse(lg.data_ptr(), ld, t.data_ptr(), row_loss.data_ptr(),
                                                               inv.data_ptr(), 0.0, rows, V, ignore_index, int(need_grad),
                                                               _native.ptr(db_pre), 0, _native.ptr(ws), nb,
                                                               lse.data_ptr(), st), "xent_rows_lse")
                    else:
                        _native.check(lib.b200ml_xent_rows(lg.data_ptr(), ld, t[r0:r1].data_ptr(),
                                                           row_loss[r0:r1].data_ptr(), inv.data_ptr(), 0.0, r1 - r0, V,
                                                           ignore_index, int(need_grad), _native.ptr(db_pre), int(r0 > 0),
                                                           _native.ptr(ws), nb, st), "xent_rows")
                else:
                    _native.check(lib.b200ml_xent_fwd_bwd(lg.data_ptr(), 0, ld, t[r0:r1].data_ptr(),
                                                          row_loss[r0:r1].data_ptr(), inv.data_ptr(), 0.0, r1 - r0, V,
                                                          ignore_index, int(need_grad), st), "xent_fwd_bwd")
                if dh is not None:
                    gemm(lg[:, :V], w16, b_mn=True, out=dh[r0:r1])  # d(loss)/dh for an upstream gradient of 1
            _native.check(lib.b200ml_loss_reduce(row_loss.data_ptr(), rows, inv.data_ptr(), 0.0, loss.data_ptr(), st),
                          "loss_reduce")
            ctx.save_for_backward(h2, logits, dh if dh is not None else h2.new_empty(0),
                                  db_pre if db_pre is not None else h2.new_empty(0))
            ctx.V = V
            ctx.has_dh = dh is not None
            ctx.has_db = db_pre is not None
            return loss[0]
        zf = h2.float() @ weight.float().t()
        if bias is not None:
            zf = zf + bias.float()
        valid = t != ignore_index
        cnt = valid.sum().clamp(min=1)
        lse = torch.logsumexp(zf, -1)
        picked = zf.gather(1, t.clamp(0, V - 1).unsqueeze(1)).squeeze(1)
        loss = ((lse - picked) * valid).sum() / cnt
        g = (torch.softmax(zf, -1) - torch.nn.functional.one_hot(t.clamp(0, V - 1), V)) * valid.unsqueeze(1) / cnt
        ctx.save_for_backward(h2, g, h2.new_empty(0), h2.new_empty(0))
        ctx.V = V
        ctx.has_dh = False
        ctx.has_db = False
        return loss

    @staticmethod
    def backward(ctx, gout):
        h2, g, dh_pre, db_pre = ctx.saved_tensors
        weight, bias = ctx.params
        V = ctx.V
        d = weight.shape[1]
        if _native.use_native(h2):
            gdev = gout.reshape(1).float().contiguous()
            gv = g[:, :V]
            dh = None
            if ctx.needs_input_grad[0]:
                if ctx.has_dh:   # computed chunk-wise in forward while the gradient chunk was L2-resident
                    dh = (dh_pre * gdev.to(dh_pre.dtype)).reshape(ctx.h_shape).to(ctx.h_dtype)
                else:
                    dh = gemm(gv, working_copy(weight), b_mn=True, alpha_dev=gdev)
                    dh = dh.reshape(ctx.h_shape).to(ctx.h_dtype)
            dw = db = None
            if weight.requires_grad:
                dw = deliver_grad(weight, lambda out, acc: gemm(gv, h2, a_mn=True, b_mn=True, out=out, accumulate=acc,
                                                                 alpha_dev=gdev, max_ctas=HEAD_WGRAD_CTAS), (V, d))
            if bias is not None and bias.requires_grad:
                def _db(out, acc):
                    if ctx.has_db:   # column sums of the gradient were accumulated by the loss kernel itself
                        out.add_(db_pre * gdev) if acc else torch.mul(db_pre, gdev, out=out)
                    elif acc:
                        out.add_(colsum(gv) * gdev)
                    else:
                        colsum(gv, out=out)
                        out.mul_(gdev)
                db = deliver_grad(bias, _db, (V,))
            return dh, dw, db, None, None
        gf = g.float() * gout
        dh = (gf @ weight.float()).to(ctx.h_dtype).reshape(ctx.h_shape) if ctx.needs_input_grad[0] else None
        dw = deliver_grad(weight, lambda out, acc: out.add_(gf.t() @ h2.float()) if acc else out.copy_(gf.t() @ h2.float()),
                          (V, d)) if weight.requires_grad else None
        db = None
        if bias is not None and bias.requires_grad:
            db = deliver_grad(bias, lambda out, acc: out.add_(gf.sum(0)) if acc else out.copy_(gf.sum(0)), (V,))
        return dh, dw, db, None, None


def linear_cross_entropy(h, weight, bias, targets, ignore_index: int = -100):
    """mean_{valid}( CE(h·Wᵀ + b, targets) ) as a single node (head GEMM → in-place softmax-CE gradient → dgrad/wgrad)."""
    return _LinearCrossEntropy.apply(h, weight, bias, targets, ignore_index)
