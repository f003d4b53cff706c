"""Residual + dropout + LayerNorm (csrc/kernels/layernorm.cu).

Reference: transformer.py:86-101 (LayerNormalization with ``gamma``/``beta``, biased variance, eps inside the sqrt)
applied as ``norm(dropout(sublayer_out) + residual)`` at transformer.py:130-139,209-224.
"""
from __future__ import annotations

import torch

import ctypes as C

from .. import _native, _native_ext
from ..utils import workspace
from ..utils.rng import DeviceCounters
from .gemm import deliver_grad


_vp, _i, _f, _u = C.c_void_p, C.c_int, C.c_float, C.c_uint
_native_ext.register("b200ml_layernorm_bwd_bias", [_vp] * 11 + [C.c_size_t, _i, _i, _f, _vp, _u, _i, _i, _vp])


class _LayerNormResidual(torch.autograd.Function):
    """``a_bias`` (optional) is the bias of the Linear that produced ``a``: its gradient is the column sum of d_a, which
    the backward kernel has in registers anyway; the producing op is then told to skip its own bias-gradient pass
    (``skip_last_bias_grad`` of ops.gemm.mlp_chain / linear)."""

    @staticmethod
    def forward(ctx, a, x, gamma, beta, eps, p, site, a_bias=None):
        d = gamma.numel()
        lead = a.shape
        a2 = a.reshape(-1, d)
        x2 = None if x is None else x.reshape(-1, d)
        rows = a2.shape[0]
        ctx.p, ctx.site, ctx.has_x = p, site, x is not None
        ctx.params = (gamma, beta)
        ctx.a_bias = a_bias
        ctx.in_dtype = a.dtype
        if _native.use_native(a2):
            a2 = a2.to(torch.bfloat16).contiguous()
            if x2 is not None:
                x2 = x2.to(torch.bfloat16).contiguous()
            y = torch.empty_like(a2)
            z = torch.empty_like(a2)
            mean = torch.empty(rows, dtype=torch.float32, device=a.device)
            rstd = torch.empty(rows, dtype=torch.float32, device=a.device)
            rng = DeviceCounters.get(a.device).rng.data_ptr() if p > 0 else None
            lib = _native.load()
            _native.check(lib.b200ml_layernorm_fwd(a2.data_ptr(), _native.ptr(x2), gamma.data_ptr(), beta.data_ptr(),
                                                   y.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(), rows,
                                                   d, eps, p, rng, site, _native.stream_ptr()), "layernorm_fwd")
            ctx.save_for_backward(z, mean, rstd)
            ctx.mask = None
            return y.reshape(lead)
        # PyTorch restatement
        af = a2.float()
        mask = None
        if p > 0:
            mask = (torch.rand_like(af) >= p).float() / (1.0 - p)
            af = af * mask
        z = af if x2 is None else af + x2.float()
        mean = z.mean(-1)
        var = ((z - mean[:, None]) ** 2).mean(-1)
        rstd = torch.rsqrt(var + eps)
        y = (z - mean[:, None]) * rstd[:, None] * gamma.float() + beta.float()
        ctx.save_for_backward(z, mean, rstd)
        ctx.mask = mask
        return y.to(a.dtype).reshape(lead)

    @staticmethod
    def backward(ctx, dy):
        z, mean, rstd = ctx.saved_tensors
        gamma, beta = ctx.params
        d = gamma.numel()
        rows = z.shape[0]
        dy2 = dy.reshape(-1, d)
        if _native.use_native(z):
            dy2 = dy2.to(torch.bfloat16).contiguous()
            dz = torch.empty_like(z)
            da = torch.empty_like(z) if ctx.p > 0 else None
            lib = _native.load()
            nbytes = lib.b200ml_layernorm_bwd_scratch_bytes(rows, d)
            ws = workspace.get(f"ln_bwd@{workspace.lane()}", nbytes, z.device)  # one per stream lane
            rng = DeviceCounters.get(z.device).rng.data_ptr() if ctx.p > 0 else None
            slot_g = getattr(gamma, "_b200_grad", None)
            slot_b = getattr(beta, "_b200_grad", None)
            direct = slot_g is not None and slot_b is not None
            if direct:
                dg, db = slot_g, slot_b
                acc = int(bool(getattr(gamma, "_b200_grad_accumulate", False)))
            else:
                dg = torch.empty(d, dtype=torch.float32, device=z.device)
                db = torch.empty(d, dtype=torch.float32, device=z.device)
                acc = 0
            ab = ctx.a_bias
            dab = None
            if ab is not None and ab.requires_grad:
                slot_ab = getattr(ab, "_b200_grad", None)
                if slot_ab is not None:
                    dab_t, acc_ab = slot_ab, int(bool(getattr(ab, "_b200_grad_accumulate", False)))
                else:
                    dab_t, acc_ab = torch.empty(d, dtype=torch.float32, device=z.device), 0
                _native.check(lib.b200ml_layernorm_bwd_bias(
                    dy2.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), dz.data_ptr(),
                    _native.ptr(da), dg.data_ptr(), db.data_ptr(), dab_t.data_ptr(), ws.data_ptr(), ws.numel(), rows, d,
                    ctx.p, rng, ctx.site, acc, acc_ab, _native.stream_ptr()), "layernorm_bwd_bias")
                if slot_ab is not None:
                    ready = getattr(ab, "_b200_ready", None)
                    if ready is not None:
                        ready()
                else:
                    dab = dab_t
            else:
                _native.check(lib.b200ml_layernorm_bwd(dy2.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                                       gamma.data_ptr(), dz.data_ptr(), _native.ptr(da), dg.data_ptr(),
                                                       db.data_ptr(), ws.data_ptr(), ws.numel(), rows, d, ctx.p, rng,
                                                       ctx.site, acc, _native.stream_ptr()), "layernorm_bwd")
            if direct:
                for prm in (gamma, beta):
                    ready = getattr(prm, "_b200_ready", None)
                    if ready is not None:
                        ready()
                dg = db = None
            dz_r = dz.reshape(dy.shape)
            da_r = dz_r if da is None else da.reshape(dy.shape)
            return da_r, (dz_r if ctx.has_x else None), dg, db, None, None, None, dab
        dyf = dy2.float()
        xh = (z - mean[:, None]) * rstd[:, None]
        dg = (dyf * xh).sum(0)
        db = dyf.sum(0)
        gy = dyf * gamma.float()
        s1 = gy.mean(-1, keepdim=True)
        s2 = (gy * xh).mean(-1, keepdim=True)
        dz = rstd[:, None] * (gy - s1 - xh * s2)
        da = dz if ctx.mask is None else dz * ctx.mask
        dgo = deliver_grad(gamma, lambda out, acc: out.add_(dg) if acc else out.copy_(dg), (d,))
        dbo = deliver_grad(beta, lambda out, acc: out.add_(db) if acc else out.copy_(db), (d,))
        dab = None
        if ctx.a_bias is not None and ctx.a_bias.requires_grad:
            s_a = da.sum(0)
            dab = deliver_grad(ctx.a_bias, lambda out, acc: out.add_(s_a) if acc else out.copy_(s_a), (d,))
        return (da.to(ctx.in_dtype).reshape(dy.shape), dz.to(ctx.in_dtype).reshape(dy.shape) if ctx.has_x else None,
                dgo, dbo, None, None, None, dab)


def layernorm_residual(a, residual, gamma, beta, eps: float = 1e-5, dropout_p: float = 0.0, site: int = 0,
                       training: bool = True, a_bias=None):
    """LayerNorm(dropout(a) + residual) over the last dimension (``residual`` may be None).  ``a_bias``: the bias of the
    Linear that produced ``a`` when that op was told to leave its bias gradient to this one."""
    return _LayerNormResidual.apply(a, residual, gamma, beta, eps, dropout_p if training else 0.0, site, a_bias)
