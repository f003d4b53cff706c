"""Fused flat optimizers (csrc/kernels/optim.cu): one launch updates fp32 master weights, Adam moments and the bf16
working copy.  Semantics of torch.optim.SGD(lr) (distributed_cnn.py:138; distributed_multilayer_perceptron.py:111)
and torch.optim.Adam(lr) (distributed_lstm.py:141; pytorch_machine_translator.py:129)."""
from __future__ import annotations

import torch

from .. import _native
from ..parallel.flat import FlatParams
from ..utils.rng import DeviceCounters

KINDS = {"sgd": 1, "adam": 2}


def optim_step_flat(flat_p, flat_g, m, v, w16, kind: int, lr, betas, eps, weight_decay, step_tensor, grad_scale=1.0,
                    start: int = 0, n: int | None = None):
    """Apply one update to flat[start:start+n].  ``step_tensor`` is the (already incremented) int32 device step."""
    n = flat_p.numel() - start if n is None else n
    if n <= 0:
        return
    if _native.use_native(flat_p):
        lib = _native.load()
        f4, b2 = 4 * start, 2 * start
        _native.check(lib.b200ml_optim_step(
            flat_p.data_ptr() + f4, flat_g.data_ptr() + (b2 if flat_g.dtype == torch.bfloat16 else f4),
            int(flat_g.dtype == torch.bfloat16), None if m is None else m.data_ptr() + f4,
            None if v is None else v.data_ptr() + f4, None if w16 is None else w16.data_ptr() + b2, n, kind, lr,
            betas[0], betas[1], eps, weight_decay, step_tensor.data_ptr(), grad_scale, _native.stream_ptr()),
            "optim_step")
        return
    sl = slice(start, start + n)
    g = flat_g[sl].float() * grad_scale
    p = flat_p[sl]
    if weight_decay:
        g = g + weight_decay * p
    if kind == 2:
        t = float(step_tensor.item())
        m[sl].mul_(betas[0]).add_(g, alpha=1 - betas[0])
        v[sl].mul_(betas[1]).addcmul_(g, g, value=1 - betas[1])
        denom = v[sl].sqrt() / (1 - betas[1] ** t) ** 0.5 + eps
        p.addcdiv_(m[sl], denom, value=-lr / (1 - betas[0] ** t))
    else:
        p.add_(g, alpha=-lr)
    if w16 is not None:
        w16[sl].copy_(p)


class FlatOptimizer:
    """Single-process optimizer over FlatParams (the multi-GPU engine fuses the same update into its all-reduce)."""

    def __init__(self, params, kind: str = "sgd", lr: float = 1e-2, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, bf16_copy: bool = True):
        self.flat = params if isinstance(params, FlatParams) else FlatParams(list(params), bf16_copy=bf16_copy)
        self.kind = KINDS[kind]
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        dev = self.flat.device
        self.m = torch.zeros_like(self.flat.flat_p) if self.kind == 2 else None
        self.v = torch.zeros_like(self.flat.flat_p) if self.kind == 2 else None
        self.counters = DeviceCounters.get(dev)
        self.step_tensor = torch.zeros(1, dtype=torch.int32, device=dev)

    @property
    def param_groups(self):
        return [{"params": self.flat.params, "lr": self.lr}]

    def zero_grad(self, set_to_none: bool = True) -> None:
        self.flat.zero_accumulating_slots()
        for p in self.flat.params:
            p.grad = None

    def step(self, grad_scale: float = 1.0) -> None:
        for p in self.flat.params:  # gradients assigned by hand (not through autograd) never hit the hook
            if p.grad is not None:
                p._b200_grad.copy_(p.grad.reshape(-1))
                p.grad = None
        # one one-thread kernel bumps this optimizer's step AND the dropout RNG step: backward has consumed this step's
        # masks by now, and without the bump every training step would draw the identical masks
        self.counters.advance(rng=True, opt_tensor=self.step_tensor)
        optim_step_flat(self.flat.flat_p, self.flat.flat_g, self.m, self.v, self.flat.flat_w16, self.kind, self.lr,
                        self.betas, self.eps, self.weight_decay, self.step_tensor, grad_scale)

    def state_dict(self):
        return {"kind": self.kind, "lr": self.lr, "betas": self.betas, "eps": self.eps, "step": int(self.step_tensor.item()),
                "m": None if self.m is None else self.m.clone(), "v": None if self.v is None else self.v.clone()}

    def load_state_dict(self, sd) -> None:
        self.step_tensor.fill_(sd["step"])
        if self.m is not None and sd.get("m") is not None:
            self.m.copy_(sd["m"])
            self.v.copy_(sd["v"])
