"""Stacked LSTM on the GPU (csrc/kernels/lstm.cu + the tcgen05 GEMM).

Per layer, forward:   x16 = dropout(x).bf16  ->  pre = x16·W_ihᵀ + (b_ih + b_hh)   [one GEMM over all B·T rows]
                      ->  persistent recurrence kernel (h_t = cell(pre_t + h_{t-1}·W_hhᵀ), fp32)
           backward:  persistent reverse recurrence -> dG [B·T, 128] (bf16)
                      ->  dW_ih = dGᵀ·x16,  dW_hh = dGᵀ·H_prev,  db = colsum(dG),  dX = dG·W_ih   [hoisted GEMMs]
Reference: nn.LSTM(E, 32, num_layers=2, batch_first=True, dropout=0.5) (distributed_lstm.py:118-119,131).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .. import _native, _native_ext
from ..utils.rng import DeviceCounters
from .gemm import colsum, deliver_grad, gemm

vp, i, f, ll, u = C.c_void_p, C.c_int, C.c_float, C.c_longlong, C.c_uint
_native_ext.register("b200ml_lstm_rec_fwd", [vp] * 9 + [i, i, i, vp])
_native_ext.register("b200ml_lstm_rec_bwd", [vp] * 10 + [i, i, i, vp])
_native_ext.register("b200ml_dropout_cast", [vp, vp, ll, f, vp, u, vp])
_native_ext.register("b200ml_dropout_bwd", [vp, i, vp, ll, f, vp, u, vp])
_native_ext.register("b200ml_lstm2_fwd", [vp] * 16 + [i, i, i, f, vp, u, vp])
_native_ext.register("b200ml_lstm2_bwd", [vp] * 15 + [i, i, i, f, vp, u, vp])
# two-layer wavefront kernels (csrc/kernels/lstm_wave.cu): hidden 32, two layers - the reference configuration
WAVEFRONT = os.environ.get("B200ML_LSTM_WAVE", "1") != "0"


def _param_grads(grads, layer, weights, dg, x16, h0_l, h_out, B, T, H, K):
    """Hoisted parameter gradients of one layer from its gate gradients dG [B*T, 4H] (bf16)."""
    w_ih, w_hh, b_ih, b_hh = weights[4 * layer:4 * layer + 4]
    hprev16 = torch.cat([h0_l.unsqueeze(1), h_out[:, :-1, :]], dim=1).reshape(B * T, H).to(torch.bfloat16)
    if w_ih.requires_grad:
        grads[4 * layer] = deliver_grad(w_ih, lambda out, acc, g=dg, a=x16: gemm(
            g, a, a_mn=True, b_mn=True, out=out, accumulate=acc), (4 * H, K))
    if w_hh.requires_grad:
        grads[4 * layer + 1] = deliver_grad(w_hh, lambda out, acc, g=dg, a=hprev16: gemm(
            g, a, a_mn=True, b_mn=True, out=out, accumulate=acc), (4 * H, H))
    for j, b in ((2, b_ih), (3, b_hh)):
        if b.requires_grad:
            grads[4 * layer + j] = deliver_grad(b, lambda out, acc, g=dg: colsum(g, out=out, accumulate=acc), (4 * H,))


class _LSTMStack(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, h0, c0, num_layers, dropout_p, site, *weights):
        lib = _native.load()
        dev = x.device
        B, T, _ = x.shape
        H = weights[1].shape[1]
        st = _native.stream_ptr()
        x = x.float().contiguous()
        h0 = h0.float().contiguous()
        c0 = c0.float().contiguous()
        rng = DeviceCounters.get(dev).rng.data_ptr() if dropout_p > 0 else None
        need_grad = any(w.requires_grad for w in weights) or x.requires_grad
        inp = x
        saved, meta = [], []
        h_n = torch.empty(num_layers, B, H, dtype=torch.float32, device=dev)
        c_n = torch.empty(num_layers, B, H, dtype=torch.float32, device=dev)
        ctx.fused = WAVEFRONT and num_layers == 2 and H == 32 and weights[4].shape[1] == 32
        if ctx.fused:
            # both layers in one persistent kernel, layer 1 one step behind layer 0; only layer 0's input projection is
            # a hoisted GEMM, the inter-layer dropout and layer 1's input projection happen inside the kernel
            w_ih0, w_hh0, b_ih0, b_hh0, w_ih1, w_hh1, b_ih1, b_hh1 = weights
            K0 = x.shape[2]
            x16 = torch.empty(B * T, K0, dtype=torch.bfloat16, device=dev)
            _native.check(lib.b200ml_dropout_cast(x.data_ptr(), x16.data_ptr(), B * T * K0, 0.0, None, site, st),
                          "dropout_cast")
            w_ih0_16 = w_ih0.detach().to(torch.bfloat16)
            pre0 = gemm(x16, w_ih0_16, bias=(b_ih0.detach().float() + b_hh0.detach().float()).contiguous(),
                        out_dtype=torch.float32)
            bias1 = (b_ih1.detach().float() + b_hh1.detach().float()).contiguous()
            w_ih1_32 = w_ih1.detach().float().contiguous()
            w_hh0_32, w_hh1_32 = w_hh0.detach().float().contiguous(), w_hh1.detach().float().contiguous()
            f32 = dict(dtype=torch.float32, device=dev)
            h_out0, h_out1 = torch.empty(B, T, H, **f32), torch.empty(B, T, H, **f32)
            gates0 = torch.empty(B, T, 4 * H, **f32) if need_grad else None
            gates1 = torch.empty(B, T, 4 * H, **f32) if need_grad else None
            c_seq0 = torch.empty(B, T, H, **f32) if need_grad else None
            c_seq1 = torch.empty(B, T, H, **f32) if need_grad else None
            x16_1 = torch.empty(B * T, H, dtype=torch.bfloat16, device=dev)
            _native.check(lib.b200ml_lstm2_fwd(
                pre0.data_ptr(), w_hh0_32.data_ptr(), w_ih1_32.data_ptr(), bias1.data_ptr(), w_hh1_32.data_ptr(),
                h0.data_ptr(), c0.data_ptr(), h_out0.data_ptr(), _native.ptr(gates0), _native.ptr(c_seq0),
                h_out1.data_ptr(), _native.ptr(gates1), _native.ptr(c_seq1), x16_1.data_ptr(), h_n.data_ptr(),
                c_n.data_ptr(), B, T, H, dropout_p, rng, site + 1, st), "lstm2_fwd")
            ctx.save_for_backward(x16, w_ih0_16, w_hh0_32, h_out0, gates0, c_seq0,
                                  x16_1, w_ih1_32, w_hh1_32, h_out1, gates1, c_seq1)
            ctx.meta = [(0.0, K0), (dropout_p, H)]
            ctx.state = (h0, c0, num_layers, site, B, T, H)
            ctx.weights = weights
            return h_out1, h_n, c_n
        for layer in range(num_layers):
            w_ih, w_hh, b_ih, b_hh = weights[4 * layer:4 * layer + 4]
            K = inp.shape[2]
            p = dropout_p if layer > 0 else 0.0
            x16 = torch.empty(B * T, K, dtype=torch.bfloat16, device=dev)
            _native.check(lib.b200ml_dropout_cast(inp.data_ptr(), x16.data_ptr(), B * T * K, p, rng if p > 0 else None,
                                                  site + layer, st), "dropout_cast")
            w_ih16 = w_ih.detach().to(torch.bfloat16)
            bias = (b_ih.detach().float() + b_hh.detach().float()).contiguous()
            pre = gemm(x16, w_ih16, bias=bias, out_dtype=torch.float32)          # [B*T, 4H]
            w_hh32 = w_hh.detach().float().contiguous()
            h_out = torch.empty(B, T, H, dtype=torch.float32, device=dev)
            gates = torch.empty(B, T, 4 * H, dtype=torch.float32, device=dev) if need_grad else None
            c_seq = torch.empty(B, T, H, dtype=torch.float32, device=dev) if need_grad else None
            _native.check(lib.b200ml_lstm_rec_fwd(pre.data_ptr(), w_hh32.data_ptr(), h0[layer].data_ptr(),
                                                  c0[layer].data_ptr(), h_out.data_ptr(), _native.ptr(gates),
                                                  _native.ptr(c_seq), h_n[layer].data_ptr(), c_n[layer].data_ptr(), B,
                                                  T, H, st), "lstm_rec_fwd")
            saved += [x16, w_ih16, w_hh32, h_out, gates, c_seq]
            meta.append((p, K))
            inp = h_out
        # save_for_backward (not attributes): the last h_out is also an output of this node, and an attribute reference
        # to it would form a cycle node -> output -> grad_fn that only the garbage collector can break
        ctx.save_for_backward(*saved)
        ctx.meta = meta
        ctx.state = (h0, c0, num_layers, site, B, T, H)
        ctx.weights = weights
        return inp, h_n, c_n

    @staticmethod
    def backward(ctx, d_out, d_hn, d_cn):
        lib = _native.load()
        h0, c0, num_layers, site, B, T, H = ctx.state
        weights = ctx.weights
        saved = ctx.saved_tensors
        dev = d_out.device
        st = _native.stream_ptr()
        rngp = DeviceCounters.get(dev).rng.data_ptr()
        d_seq = d_out.float().contiguous()
        d_hn = None if d_hn is None else d_hn.float().contiguous()
        d_cn = None if d_cn is None else d_cn.float().contiguous()
        dh0 = torch.empty(num_layers, B, H, dtype=torch.float32, device=dev)
        dc0 = torch.empty(num_layers, B, H, dtype=torch.float32, device=dev)
        grads = [None] * len(weights)
        dx = None
        if ctx.fused:
            x16, w_ih0_16, w_hh0_32, h_out0, gates0, c_seq0, x16_1, w_ih1_32, w_hh1_32, h_out1, gates1, c_seq1 = saved
            p1, K0 = ctx.meta[1][0], ctx.meta[0][1]
            dg0 = torch.empty(B * T, 4 * H, dtype=torch.bfloat16, device=dev)
            dg1 = torch.empty(B * T, 4 * H, dtype=torch.bfloat16, device=dev)
            _native.check(lib.b200ml_lstm2_bwd(
                w_hh0_32.data_ptr(), w_ih1_32.data_ptr(), w_hh1_32.data_ptr(), c0.data_ptr(), gates0.data_ptr(),
                c_seq0.data_ptr(), gates1.data_ptr(), c_seq1.data_ptr(), _native.ptr(d_seq), _native.ptr(d_hn),
                _native.ptr(d_cn), dg0.data_ptr(), dg1.data_ptr(), dh0.data_ptr(), dc0.data_ptr(), B, T, H, p1,
                rngp if p1 > 0 else None, site + 1, st), "lstm2_bwd")
            # input gradient first (reads W_ih0), then the parameter gradients (may release their bucket)
            if ctx.needs_input_grad[0]:
                dx = gemm(dg0, w_ih0_16, b_mn=True, out_dtype=torch.float32).view(B, T, K0)
            _param_grads(grads, 1, weights, dg1, x16_1, h0[1], h_out1, B, T, H, H)
            _param_grads(grads, 0, weights, dg0, x16, h0[0], h_out0, B, T, H, K0)
            return (dx, dh0, dc0, None, None, None, *grads)
        for layer in range(num_layers - 1, -1, -1):
            x16, w_ih16, w_hh32, h_out, gates, c_seq = saved[6 * layer:6 * layer + 6]
            p, K = ctx.meta[layer]
            w_ih, w_hh, b_ih, b_hh = weights[4 * layer:4 * layer + 4]
            dg = torch.empty(B * T, 4 * H, dtype=torch.bfloat16, device=dev)
            _native.check(lib.b200ml_lstm_rec_bwd(
                w_hh32.data_ptr(), c0[layer].data_ptr(), gates.data_ptr(), c_seq.data_ptr(), _native.ptr(d_seq),
                None if d_hn is None else d_hn[layer].data_ptr(), None if d_cn is None else d_cn[layer].data_ptr(),
                dg.data_ptr(), dh0[layer].data_ptr(), dc0[layer].data_ptr(), B, T, H, st), "lstm_rec_bwd")
            # input gradient first (reads W_ih), then the parameter gradients (may release their bucket)
            need_dx = layer > 0 or ctx.needs_input_grad[0]
            if need_dx:
                dxd = gemm(dg, w_ih16, b_mn=True, out_dtype=torch.float32)       # d(dropped input) [B*T, K]
                if p > 0:
                    dx = torch.empty(B, T, K, dtype=torch.float32, device=dev)
                    _native.check(lib.b200ml_dropout_bwd(dxd.data_ptr(), 1, dx.data_ptr(), B * T * K, p, rngp,
                                                         site + layer, st), "dropout_bwd")
                else:
                    dx = dxd.view(B, T, K)
            else:
                dx = None
            hprev16 = torch.cat([h0[layer].unsqueeze(1), h_out[:, :-1, :]], dim=1).reshape(B * T, H).to(torch.bfloat16)
            if w_ih.requires_grad:
                grads[4 * layer] = deliver_grad(w_ih, lambda out, acc, g=dg, a=x16: gemm(
                    g, a, a_mn=True, b_mn=True, out=out, accumulate=acc), (4 * H, K))
            if w_hh.requires_grad:
                grads[4 * layer + 1] = deliver_grad(w_hh, lambda out, acc, g=dg, a=hprev16: gemm(
                    g, a, a_mn=True, b_mn=True, out=out, accumulate=acc), (4 * H, H))
            for j, b in ((2, b_ih), (3, b_hh)):
                if b.requires_grad:
                    grads[4 * layer + j] = deliver_grad(b, lambda out, acc, g=dg: colsum(g, out=out, accumulate=acc),
                                                        (4 * H,))
            d_seq = dx
        return (dx if ctx.needs_input_grad[0] else None, dh0, dc0, None, None, None, *grads)


def lstm_forward(x, h0, c0, weights, num_layers, dropout_p, site):
    assert weights[1].shape[1] in (32, 64, 128), "the LSTM recurrence kernels cover hidden_size 32 (the reference value), 64 and 128"
    return _LSTMStack.apply(x, h0, c0, num_layers, dropout_p, site, *weights)
