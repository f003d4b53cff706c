"""TinyVGG of the reference (``FashionMNISTModel``, distributed_cnn.py:47-86 / pytorch_cnn.py:12-49):
block_1 = conv3x3-ReLU-conv3x3-ReLU-maxpool2, block_2 = the same, classifier = Flatten + Linear.

Same constructor and state_dict keys (block_{1,2}.{0,2}.{weight,bias}, classifier.1.{weight,bias}); the reference
hard-codes ``hidden_units*7*7`` (28×28 inputs only) — here the classifier width follows ``image_size`` so the
BASELINE.json CIFAR-shape preset (3×32×32) works, with 28 as the default.  Each conv+bias+ReLU(+2×2 max-pool) is one
implicit-GEMM kernel (ops/conv.py).
"""
from __future__ import annotations

import torch
from torch import nn

from ..ops import conv as _conv
from ..ops.gemm import linear as _linear


class FashionMNISTModel(nn.Module):
    def __init__(self, input_shape: int, hidden_units: int, output_shape: int, image_size=28):
        super().__init__()
        hw = (image_size, image_size) if isinstance(image_size, int) else tuple(image_size)
        assert hw[0] % 4 == 0 and hw[1] % 4 == 0, "two 2x2 max-pools need H, W divisible by 4"
        self.image_size = hw
        self.block_1 = nn.Sequential(
            nn.Conv2d(in_channels=input_shape, out_channels=hidden_units, kernel_size=3, stride=1, padding=1),
            nn.ReLU(),
            nn.Conv2d(in_channels=hidden_units, out_channels=hidden_units, kernel_size=3, stride=1, padding=1),
            nn.ReLU(),
            nn.MaxPool2d(kernel_size=2, stride=2))
        self.block_2 = nn.Sequential(
            nn.Conv2d(hidden_units, hidden_units, 3, padding=1),
            nn.ReLU(),
            nn.Conv2d(hidden_units, hidden_units, 3, padding=1),
            nn.ReLU(),
            nn.MaxPool2d(2))
        self.classifier = nn.Sequential(
            nn.Flatten(),
            nn.Linear(in_features=hidden_units * (hw[0] // 4) * (hw[1] // 4), out_features=output_shape))

    def forward(self, x: torch.Tensor):
        for blk in (self.block_1, self.block_2):
            x = _conv.conv3x3_relu(x, blk[0].weight, blk[0].bias, pool=False)
            x = _conv.conv3x3_relu(x, blk[2].weight, blk[2].bias, pool=True)
        x = x.flatten(1)
        fc = self.classifier[1]
        return _linear(x, fc.weight, fc.bias).float()
