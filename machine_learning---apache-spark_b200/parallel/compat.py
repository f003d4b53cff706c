"""The reference's training-loop vocabulary on top of the engine: ``DistributedDataParallel(model)`` plus
``SGD(model.parameters(), lr)`` / ``Adam(model.parameters(), lr)`` with ``zero_grad()`` / ``step()``.

The reference writes (distributed_cnn.py:138,156,175-183; distributed_lstm.py:141,160,186-194;
distributed_multilayer_perceptron.py:106,111,130-138)

    DDP_model = DDP(cnn_model)
    optimizer = torch.optim.SGD(cnn_model.parameters(), lr=0.01)
    ...
    predictions = cnn_model(data)            # (or DDP_model(data))
    loss = loss_fn(predictions, target)
    optimizer.zero_grad(); loss.backward(); optimizer.step()

and a user who switches keeps exactly that shape:

    from b200ml.parallel import DistributedDataParallel as DDP
    from b200ml.optim import SGD
    DDP_model = DDP(cnn_model)
    optimizer = SGD(cnn_model.parameters(), lr=0.01)

Underneath there is ONE object, the ``DataParallelEngine`` (flat buffers, bucket reducer, fused all-reduce + optimizer
kernel): gradient reduction and the weight update are one kernel here, so the wrapper and the optimizer are two faces
of it.  The wrapper tags the parameters; the optimizer finds the wrapper through them (the reference builds its
optimizer on the BARE model's parameters - the same tensors) and creates the engine with its own hyper-parameters.
Without a wrapper (the single-process scripts: pytorch_cnn.py:105, pytorch_lstm.py:124) the optimizer drives a
``FlatOptimizer``.  ``zero_grad()`` may sit before the forward pass or between forward and backward, as in the
reference; a forward through the wrapper opens the step by itself if nothing has yet.
"""
from __future__ import annotations

import weakref

import torch
from torch import nn

from ..ops.optim import FlatOptimizer
from .engine import DataParallelEngine


class DistributedDataParallel(nn.Module):
    """``torch.nn.parallel.DistributedDataParallel`` as the reference uses it: wraps the module (``state_dict`` keys get
    the ``module.`` prefix, like torch's), rank 0's parameters are broadcast and the cross-rank layout is verified when
    the engine is created, gradients are averaged over the ranks during backward.  ``device_ids`` / ``output_device`` /
    ``bucket_cap_mb`` are accepted with torch's meaning; anything else torch's constructor takes is ignored."""

    def __init__(self, module: nn.Module, device_ids=None, output_device=None, bucket_cap_mb: float = 25.0,
                 allreduce: str = "auto", sync_gradients: bool = True, **_ignored):
        super().__init__()
        self.module = module
        self._cfg = dict(bucket_mb=float(bucket_cap_mb), allreduce=allreduce, sync_gradients=sync_gradients)
        self._engine: DataParallelEngine | None = None
        self._open = False                      # a step has been begun and not finished
        ref = weakref.ref(self)
        for p in module.parameters():
            p._b200_ddp = ref

    # ------------------------------------------------------------------------------------------------ engine binding
    def _attach(self, kind: str, lr: float, betas, eps: float, weight_decay: float) -> DataParallelEngine:
        if self._engine is not None:
            raise RuntimeError("this DistributedDataParallel module already has an optimizer: gradient reduction and the "
                               "weight update are one fused kernel here, so there is one optimizer per wrapped module")
        self._engine = DataParallelEngine(self.module, optimizer=kind, lr=lr, betas=betas, eps=eps,
                                          weight_decay=weight_decay, **self._cfg)
        return self._engine

    @property
    def engine(self) -> DataParallelEngine | None:
        return self._engine

    def _begin(self) -> None:
        if self._engine is not None and not self._open:
            self._engine.begin_step()
            self._open = True

    def _finish(self) -> None:
        if self._engine is None:
            return
        if not self._open:                      # step() without zero_grad() / forward: still a (gradient-less) step
            self._engine.begin_step()
        self._engine.finish_step()
        self._open = False

    # ------------------------------------------------------------------------------------------------ nn.Module face
    def forward(self, *args, **kwargs):
        if self.training and torch.is_grad_enabled():
            self._begin()
        return self.module(*args, **kwargs)

    def close(self) -> None:
        """Release peer mappings / graphs (the scripts call it before returning the model from a worker)."""
        if self._engine is not None:
            self._engine.close()
            self._engine = None
        for p in self.module.parameters():
            if getattr(p, "_b200_ddp", None) is not None:
                p._b200_ddp = None


class _Optimizer:
    """Common part of SGD / Adam below: ``zero_grad`` / ``step`` / ``param_groups`` / ``state_dict``."""

    def __init__(self, params, kind: str, lr: float, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
        params = list(params)
        if not params:
            raise ValueError("optimizer got an empty parameter list")
        if isinstance(params[0], dict):
            raise ValueError("parameter groups are not supported: one learning rate per optimizer (as in the reference)")
        owners = {id(o): o for o in (getattr(p, "_b200_ddp", None) for p in params) if o is not None}
        wrappers = [o() for o in owners.values() if o() is not None]
        self._ddp: DistributedDataParallel | None = None
        self._flat: FlatOptimizer | None = None
        if wrappers:
            if len(wrappers) > 1:
                raise ValueError("the parameters belong to several DistributedDataParallel modules")
            ddp = wrappers[0]
            if {id(p) for p in params} != {id(p) for p in ddp.module.parameters()}:
                raise ValueError("pass ALL parameters of the wrapped module (model.parameters()): the engine updates the "
                                 "whole flat buffer in its fused all-reduce")
            ddp._attach(kind, lr, betas, eps, weight_decay)
            self._ddp = ddp
        else:
            self._flat = FlatOptimizer(params, kind=kind, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        self.defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        self.param_groups = [dict(self.defaults, params=params)]

    def zero_grad(self, set_to_none: bool = True) -> None:
        if self._ddp is not None:
            self._ddp._begin()                  # idempotent inside a step: before or after the forward pass
        else:
            self._flat.zero_grad()

    def step(self, closure=None):
        if closure is not None:
            raise ValueError("closures are not supported")
        if self._ddp is not None:
            self._ddp._finish()
        else:
            self._flat.step()

    def state_dict(self):
        return self._ddp.engine.state_dict() if self._ddp is not None else self._flat.state_dict()

    def load_state_dict(self, sd) -> None:
        (self._ddp.engine if self._ddp is not None else self._flat).load_state_dict(sd)


class SGD(_Optimizer):
    """``torch.optim.SGD(params, lr)`` (distributed_cnn.py:138, distributed_multilayer_perceptron.py:111): plain SGD."""

    def __init__(self, params, lr: float = 1e-3, momentum: float = 0.0, weight_decay: float = 0.0, **_ignored):
        if momentum:
            raise ValueError("momentum is not implemented (the reference uses plain SGD)")
        super().__init__(params, "sgd", lr, weight_decay=weight_decay)


class Adam(_Optimizer):
    """``torch.optim.Adam(params, lr)`` (distributed_lstm.py:141, pytorch_machine_translator.py:129)."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                 **_ignored):
        super().__init__(params, "adam", lr, betas=betas, eps=eps, weight_decay=weight_decay)
