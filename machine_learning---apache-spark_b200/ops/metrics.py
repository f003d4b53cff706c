"""Evaluation metrics on the device (csrc/kernels/optim.cu: argmax_accuracy_kernel).

The reference's eval loops compute ``torch.argmax(torch.softmax(predictions, dim=1), dim=1)`` and compare with the
labels on the host every batch (distributed_cnn.py:144-147,211-212; pytorch_lstm.py:73; the MLP scripts' accuracy_fn).
Softmax is monotonic, so the arg-max of the logits is the same class; one kernel writes the predicted classes and adds
the number of correct rows into a device counter, and the host reads that counter once per evaluation instead of
synchronising every batch.
"""
from __future__ import annotations

import torch

from .. import _native


def argmax_accuracy(logits: torch.Tensor, targets: torch.Tensor | None = None, correct: torch.Tensor | None = None):
    """Returns ``(pred, correct)``: ``pred`` int64 [rows] = arg-max over the last dimension of ``logits`` [rows, C];
    ``correct`` is a one-element int32 device counter to which the number of rows with ``pred == targets`` has been
    ADDED (pass the tensor returned by an earlier call to accumulate over batches; ``None`` starts a new one at zero).
    """
    assert logits.dim() == 2
    rows, C = logits.shape
    if correct is None:
        correct = torch.zeros(1, dtype=torch.int32, device=logits.device)
    if _native.use_native(logits):
        z = logits.detach()
        if z.dtype not in (torch.float32, torch.bfloat16):
            z = z.float()
        if z.stride(1) != 1:
            z = z.contiguous()
        t = None if targets is None else targets.detach().reshape(-1).to(torch.int64).contiguous()
        pred = torch.empty(rows, dtype=torch.int32, device=z.device)
        _native.check(_native.load().b200ml_argmax_accuracy(
            z.data_ptr(), int(z.dtype == torch.float32), z.stride(0), _native.ptr(t), rows, C, pred.data_ptr(),
            correct.data_ptr() if t is not None else None, _native.stream_ptr()), "argmax_accuracy")
        return pred.long(), correct
    pred = logits.detach().float().argmax(1)
    if targets is not None:
        correct += (pred == targets.reshape(-1)).sum().to(torch.int32)
    return pred, correct
