"""MLP of the reference (distributed_multilayer_perceptron.py:44-53, pytorch_multilayer_perceptron.py:33-42):
layer_1 Linear(4,5) → sigmoid → layer_2 Linear(5,4) → sigmoid → layer_3 Linear(4,3).

``Multilayer_perceptor()`` with no arguments is reference-exact (state_dict keys layer_{1,2,3}.{weight,bias}); ``layers``
widens it for the BASELINE.json preset (784-dim MNIST-shape input).  The three GEMMs and both sigmoids run as one fused
chain: sigmoid in the forward epilogue, y·(1−y) in the dgrad epilogue (tcgen05 when the shapes are TMA-describable,
the SIMT kernel for the 4-5-4-3 default).
"""
from __future__ import annotations

from torch import nn

from ..ops.gemm import mlp_chain


class Multilayer_perceptor(nn.Module):
    def __init__(self, layers=(4, 5, 4, 3)):
        super().__init__()
        self.layer_sizes = tuple(layers)
        for i in range(len(layers) - 1):
            setattr(self, f"layer_{i + 1}", nn.Linear(in_features=layers[i], out_features=layers[i + 1]))
        self.sigmoid = nn.Sigmoid()

    @property
    def linears(self):
        return [getattr(self, f"layer_{i + 1}") for i in range(len(self.layer_sizes) - 1)]

    def forward(self, x):
        ls = self.linears
        acts = ["sigmoid"] * (len(ls) - 1) + [None]
        out = mlp_chain(x, [l.weight for l in ls], [l.bias for l in ls], acts)
        return out.float() if out.dtype != x.dtype and x.is_floating_point() else out


def mnist_mlp(hidden=(512, 256), classes: int = 10) -> Multilayer_perceptor:
    """BASELINE.json config 1: synthetic 784-dim MNIST-shape batches."""
    return Multilayer_perceptor((784, *hidden, classes))
