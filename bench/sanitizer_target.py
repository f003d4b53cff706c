"""Small end-to-end workload for compute-sanitizer (memcheck / racecheck / synccheck): one tiny training step of each
model family plus the stand-alone kernels that the tiny steps do not reach (CTA-pair GEMM, fused attention backward at
two tiles per side, row-buffered loss kernel, L-BFGS direction).  Sizes are minimal - sanitizer runs 10-50x slower.

    compute-sanitizer --tool memcheck python bench/sanitizer_target.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402
from b200ml.models import LSTM, FashionMNISTModel, mnist_mlp  # noqa: E402
from b200ml.ops.attention import attention  # noqa: E402
from b200ml.ops.gemm import gemm  # noqa: E402
from b200ml.ops.loss import cross_entropy, linear_cross_entropy  # noqa: E402
from b200ml.parallel.engine import DataParallelEngine  # noqa: E402
from b200ml.spark.ml import _native_direction  # noqa: E402


def step(model, kind, lr, fwd):
    eng = DataParallelEngine(model, optimizer=kind, lr=lr)
    for _ in range(2):
        eng.begin_step()
        fwd().backward()
        eng.finish_step()
    torch.cuda.synchronize()
    eng.close()


def main():
    dev = torch.device("cuda")
    torch.manual_seed(0)
    entry.smoke()                                                       # tiny Transformer-MT steps
    cnn = FashionMNISTModel(1, 10, 10).to(dev).train()                  # direct conv (layer 1) + implicit-GEMM convs
    x, y = torch.rand(4, 1, 28, 28, device=dev), torch.randint(0, 10, (4,), device=dev)
    step(cnn, "sgd", 0.01, lambda: cross_entropy(cnn(x), y))
    lstm = LSTM(500, 32, 32, 4).to(dev).train()
    tok, lab, hz = torch.randint(1, 500, (4, 17), device=dev), torch.randint(0, 4, (4,), device=dev), torch.zeros(2, 4, 32, device=dev)
    step(lstm, "adam", 1e-3, lambda: cross_entropy(lstm(tok, hz, hz)[0][:, -1, :], lab))
    mlp = mnist_mlp().to(dev).train()
    xm, ym = torch.randn(8, 784, device=dev), torch.randint(0, 10, (8,), device=dev)
    step(mlp, "sgd", 0.03, lambda: cross_entropy(mlp(xm), ym))
    # CTA-pair GEMM incl. ragged edges, and the tail split
    a = torch.randn(700, 200, device=dev).to(torch.bfloat16)
    b = torch.randn(300, 200, device=dev).to(torch.bfloat16)
    gemm(a, b, tile_n=512, bias=torch.randn(300, device=dev), act="relu")
    gemm(a, b, tile_n=512, out_dtype=torch.float32)
    # attention forward + fused backward with two tiles per side (S = 200) and a ragged cross-attention shape
    qkv = (torch.randn(1, 200, 2 * 192, device=dev) * 0.5).to(torch.bfloat16).requires_grad_()
    attention(qkv, num_heads=2, mask_mode="reference").float().sum().backward()
    # row-buffered loss kernel: several rows per CTA, ragged last vector, bias-gradient column sums
    h = (torch.randn(300, 64, device=dev) * 0.5).to(torch.bfloat16).requires_grad_()
    w = (torch.randn(1001, 64, device=dev) * 0.05).requires_grad_()
    bb = torch.zeros(1001, device=dev, requires_grad=True)
    t = torch.randint(0, 1001, (300,), device=dev)
    linear_cross_entropy(h, w, bb, t, ignore_index=0).backward()
    # L-BFGS direction (cooperative kernel)
    g = torch.randn(5000, dtype=torch.float64, device=dev)
    S = [torch.randn(5000, dtype=torch.float64, device=dev) * 0.1 for _ in range(4)]
    _native_direction(g, S, [s * 1.5 for s in S])
    torch.cuda.synchronize()
    print("sanitizer target done")


if __name__ == "__main__":
    main()
