"""Flat parameter storage: every parameter becomes a view into one fp32 master buffer, with matching views into a flat
fp32 gradient buffer and a flat bf16 working-copy buffer.

This is the memory layout the whole training path is built on:
  * gradient kernels (wgrad GEMMs, the embedding segment reduction, LayerNorm) write straight into their slot
    (``param._b200_grad``) and overwrite it — no copy-in, no ``zero_grad``;
  * the optimizer (and the fused all-reduce epilogue, parallel/engine.py) walks contiguous ranges — one launch updates
    fp32 masters, Adam moments and the bf16 working copies the GEMMs read (``param._b200_w16``);
  * ``state_dict()`` of the model is untouched: parameters keep their names, shapes and fp32 dtype.

The role c10d's Reducer plays for the reference's DDP(model) (distributed_cnn.py:156) starts here.
"""
from __future__ import annotations

import torch

ALIGN = 64  # elements; keeps every slot 256-byte aligned in fp32 and 128-byte aligned in bf16 (TMA needs 16)


def _round_up(n: int, a: int) -> int:
    return (n + a - 1) // a * a


class FlatParams:
    def __init__(self, params, bf16_copy: bool = True, order=None, allocator=None):
        params = [p for p in params if p.requires_grad]
        if order is not None:
            params = [params[i] for i in order]
        assert params, "no trainable parameters"
        self.params = params
        self.device = params[0].device
        self.offsets, off = [], 0
        for p in params:
            assert p.device == self.device and p.dtype == torch.float32, "master parameters must be fp32 on one device"
            self.offsets.append(off)
            off += _round_up(p.numel(), ALIGN)
        self.numel = off
        if allocator is None:  # default: ordinary torch allocations
            def allocator(name, numel, dtype):
                return torch.zeros(numel, dtype=dtype, device=self.device)
        # (a SymmetricAllocator hands out peer-mappable buffers for the fused all-reduce, parallel/comm.py)
        self.flat_p = allocator("flat_p", off, torch.float32)
        self.flat_g = allocator("flat_g", off, torch.float32)
        self.flat_w16 = allocator("flat_w16", off, torch.bfloat16) if bf16_copy else None
        self.zero_ranges = []  # slots of parameters flagged `_b200_zero_grad` (a backward that accumulates): cleared per step
        for p, o in zip(params, self.offsets):
            n = p.numel()
            self.flat_p[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.flat_p[o:o + n].view(p.shape)
            p._b200_grad = self.flat_g[o:o + n]
            if self.flat_w16 is not None:
                p._b200_w16 = self.flat_w16[o:o + n].view(p.shape)
                p._b200_w16_managed = True
            if getattr(p, "_b200_zero_grad", False):
                self.zero_ranges.append((o, n))
            # generic autograd path (ops without direct slot writes, CPU fallbacks): move .grad into the slot
            p.register_post_accumulate_grad_hook(self._make_hook(p))
        self.refresh_working_copy()
        self._ready_cb = None

    def _make_hook(self, p):
        def hook(param):
            if param.grad is not None:
                param._b200_grad.copy_(param.grad.reshape(-1))
                param.grad = None
                cb = getattr(param, "_b200_ready", None)
                if cb is not None:
                    cb()
        return hook

    def refresh_working_copy(self) -> None:
        if self.flat_w16 is not None:
            self.flat_w16.copy_(self.flat_p)

    def zero_accumulating_slots(self) -> None:
        for o, n in self.zero_ranges:
            self.flat_g[o:o + n].zero_()

    def slot(self, i: int):
        o = self.offsets[i]
        return o, self.params[i].numel()
