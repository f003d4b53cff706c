"""3×3 stride-1 pad-1 convolution + bias + ReLU (+ 2×2 max-pool) — K1/K2/K3 (csrc/kernels/conv.cu).
Reference: the four Conv2d+ReLU pairs and two MaxPool2d of FashionMNISTModel (distributed_cnn.py:51-73)."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import _native


def conv3x3_relu_reference(x, weight, bias, pool: bool):
    y = F.relu(F.conv2d(x.float(), weight.float(), None if bias is None else bias.float(), stride=1, padding=1))
    return F.max_pool2d(y, 2) if pool else y


def conv3x3_relu(x, weight, bias, pool: bool = False):
    """relu(conv2d(x, weight, bias, stride 1, padding 1)), optionally followed by max_pool2d(2).  NCHW in, NCHW out."""
    if _native.use_native(x):
        from . import conv_native

        return conv_native.conv3x3_relu(x, weight, bias, pool)
    return conv3x3_relu_reference(x, weight, bias, pool)
