"""Autograd wrapper of the persistent LSTM layer kernels (csrc/kernels/lstm.cu)."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native, _native_ext
from ..utils import workspace
from ..utils.rng import DeviceCounters

vp, i, f, sz, u = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_uint
_native_ext.register("b200ml_lstm_layer_fwd", [vp] * 12 + [i, i, i, i, f, vp, u, vp])
_native_ext.register("b200ml_lstm_bwd_scratch_bytes", [i], size_t_return=True)
_native_ext.register("b200ml_lstm_layer_bwd", [vp] * 19 + [sz, i, i, i, i, f, vp, u, i, vp])


class _LSTMStack(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, h0, c0, num_layers, dropout_p, site, *weights):
        lib = _native.load()
        dev = x.device
        B, T, _ = x.shape
        H = weights[1].shape[1]
        x = x.float().contiguous()
        h0 = h0.float().contiguous()
        c0 = c0.float().contiguous()
        rng = DeviceCounters.get(dev).rng.data_ptr() if dropout_p > 0 else None
        inp = x
        saved = []
        h_n = torch.empty(num_layers, B, H, dtype=torch.float32, device=dev)
        c_n = torch.empty(num_layers, B, H, dtype=torch.float32, device=dev)
        need_grad = any(w.requires_grad for w in weights) or x.requires_grad
        for layer in range(num_layers):
            w_ih, w_hh, b_ih, b_hh = [w.detach().float().contiguous() for w in weights[4 * layer:4 * layer + 4]]
            K = inp.shape[2]
            h_out = torch.empty(B, T, H, dtype=torch.float32, device=dev)
            gates = torch.empty(T, B, 4, H, dtype=torch.float32, device=dev) if need_grad else None
            c_seq = torch.empty(T, B, H, dtype=torch.float32, device=dev) if need_grad else None
            p = dropout_p if layer > 0 else 0.0
            _native.check(lib.b200ml_lstm_layer_fwd(
                inp.data_ptr(), w_ih.data_ptr(), w_hh.data_ptr(), b_ih.data_ptr(), b_hh.data_ptr(),
                h0[layer].data_ptr(), c0[layer].data_ptr(), h_out.data_ptr(), _native.ptr(gates), _native.ptr(c_seq),
                h_n[layer].data_ptr(), c_n[layer].data_ptr(), B, T, K, H, p, rng if p > 0 else None, site + layer,
                _native.stream_ptr()), "lstm_layer_fwd")
            saved.append((inp, w_ih, w_hh, h_out, gates, c_seq, p))
            inp = h_out
        ctx.saved = saved
        ctx.state = (h0, c0, num_layers, site, B, T, H)
        ctx.weights = weights
        ctx.mark_non_differentiable()
        return inp, h_n, c_n

    @staticmethod
    def backward(ctx, d_out, d_hn, d_cn):
        lib = _native.load()
        h0, c0, num_layers, site, B, T, H = ctx.state
        weights = ctx.weights
        dev = d_out.device
        rngp = DeviceCounters.get(dev).rng.data_ptr()
        ws = workspace.get("lstm_bwd", lib.b200ml_lstm_bwd_scratch_bytes(B), dev)
        d_seq = d_out.float().contiguous()
        d_hn = None if d_hn is None else d_hn.float().contiguous()
        d_cn = None if d_cn is None else d_cn.float().contiguous()
        dh0 = torch.empty(num_layers, B, H, dtype=torch.float32, device=dev)
        dc0 = torch.empty(num_layers, B, H, dtype=torch.float32, device=dev)
        grads = [None] * len(weights)
        dx = None
        for layer in range(num_layers - 1, -1, -1):
            inp, w_ih, w_hh, h_out, gates, c_seq, p = ctx.saved[layer]
            K = inp.shape[2]
            prm = weights[4 * layer:4 * layer + 4]
            slots = [getattr(w, "_b200_grad", None) for w in prm]
            direct = all(s is not None for s in slots)
            if direct:
                outs = slots
                acc = int(bool(getattr(prm[0], "_b200_grad_accumulate", False)))
            else:
                outs = [torch.empty(w.numel(), dtype=torch.float32, device=dev) for w in prm]
                acc = 0
            need_dx = layer > 0 or ctx.needs_input_grad[0]
            dx = torch.empty(B, T, K, dtype=torch.float32, device=dev) if need_dx else None
            _native.check(lib.b200ml_lstm_layer_bwd(
                inp.data_ptr(), w_ih.data_ptr(), w_hh.data_ptr(), h0[layer].data_ptr(), c0[layer].data_ptr(),
                h_out.data_ptr(), gates.data_ptr(), c_seq.data_ptr(), _native.ptr(d_seq),
                None if d_hn is None else d_hn[layer].data_ptr(), None if d_cn is None else d_cn[layer].data_ptr(),
                _native.ptr(dx), dh0[layer].data_ptr(), dc0[layer].data_ptr(), outs[0].data_ptr(), outs[1].data_ptr(),
                outs[2].data_ptr(), outs[3].data_ptr(), ws.data_ptr(), ws.numel(), B, T, K, H, p,
                rngp if p > 0 else None, site + layer, acc, _native.stream_ptr()), "lstm_layer_bwd")
            if direct:
                for w in prm:
                    ready = getattr(w, "_b200_ready", None)
                    if ready is not None:
                        ready()
            else:
                for j, w in enumerate(prm):
                    grads[4 * layer + j] = outs[j].view_as(w)
            d_seq = dx
        return (dx if ctx.needs_input_grad[0] else None, dh0, dc0, None, None, None, *grads)


def lstm_forward(x, h0, c0, weights, num_layers, dropout_p, site):
    return _LSTMStack.apply(x, h0, c0, num_layers, dropout_p, site, *weights)
