"""Diagnostic only (never a benchmark number): per-kernel device time of eager training steps via torch.profiler
(CUPTI activity records: warm caches, natural concurrency) — complements the serialised ncu launch lists."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("B200ML_PDL", "0")  # early-launched kernels would be charged for the time they spend waiting
os.environ.setdefault("B200ML_WGRAD_STREAM", "0")
os.environ.setdefault("B200ML_OVERLAP_STREAMS", "0")
os.environ.setdefault("B200ML_OPT_OVERLAP", "0")
from b200ml.models.transformer import Transformer  # noqa: E402
from b200ml.ops.loss import linear_cross_entropy  # noqa: E402
from b200ml.parallel.engine import DataParallelEngine  # noqa: E402

import argparse  # noqa: E402

_ap = argparse.ArgumentParser()
_ap.add_argument("--model", default="transformer", choices=["transformer", "cnn", "cnn_cifar", "lstm", "mlp"])
_args = _ap.parse_args()
dev = torch.device("cuda")
torch.manual_seed(0)
if _args.model != "transformer":
    from b200ml.models import LSTM, FashionMNISTModel, mnist_mlp
    from b200ml.ops.loss import cross_entropy

    Bs = 32
    if _args.model.startswith("cnn"):
        ch, hw = (1, 28) if _args.model == "cnn" else (3, 32)
        model = FashionMNISTModel(ch, 10, 10, image_size=hw).to(dev).train()
        eng = DataParallelEngine(model, optimizer="sgd", lr=0.01)
        x, y = torch.rand(Bs, ch, hw, hw, device=dev), torch.randint(0, 10, (Bs,), device=dev)
    elif _args.model == "lstm":
        model = LSTM(95815, 32, 32, 4).to(dev).train()
        eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)
        x, y = torch.randint(1, 95815, (Bs, 129), device=dev), torch.randint(0, 4, (Bs,), device=dev)
    else:
        model = mnist_mlp().to(dev).train()
        eng = DataParallelEngine(model, optimizer="sgd", lr=0.03)
        x, y = torch.randn(Bs, 784, device=dev), torch.randint(0, 10, (Bs,), device=dev)
    hz = torch.zeros(2, Bs, 32, device=dev)

    def small_step():
        eng.begin_step()
        if _args.model == "lstm":
            pred, _, _ = model(x, hz, hz)
            loss = cross_entropy(pred[:, -1, :], y)
        else:
            loss = cross_entropy(model(x), y)
        loss.backward()
        eng.finish_step()

    for _ in range(3):
        small_step()
    torch.cuda.synchronize()
    N = 10
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(N):
            small_step()
        torch.cuda.synchronize()
    rows = [(e.key, e.device_time_total / N, e.count / N) for e in prof.key_averages() if e.device_time_total > 0]
    rows.sort(key=lambda r: -r[1])
    print(f"[{_args.model}] sum of kernel time per step: {sum(r[1] for r in rows):.0f} us")
    for k, t, c in rows[:40]:
        print(f"{t:9.1f} us  x{c:5.1f}  {k[:100]}")
    sys.exit(0)
B, S = 32, 200
model = Transformer(512, 1024, 8, 0.1, 1, S, 19214, 10837, 19214).to(dev).train()
eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)
src = torch.randint(1, 10837, (B, S), device=dev)
trg = torch.randint(1, 19214, (B, S), device=dev)
la = (torch.tril(torch.ones(S, S, device=dev)) == 0)[None, None]


def step():
    eng.begin_step()
    h = model(src, trg, (src != 0)[:, None, None, :], la, la, return_hidden=True)
    loss = linear_cross_entropy(h, model.linear.weight, model.linear.bias, trg, ignore_index=0)
    loss.backward()
    eng.finish_step()


for _ in range(3):
    step()
torch.cuda.synchronize()
N = 5
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(N):
        step()
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / N, e.count / N) for e in prof.key_averages() if e.device_time_total > 0]
rows.sort(key=lambda r: -r[1])
tot = sum(r[1] for r in rows)
print(f"sum of kernel time per step: {tot:.0f} us")
for k, t, c in rows[:40]:
    print(f"{t:9.1f} us  x{c:5.1f}  {k[:100]}")
