"""Profiling target for ncu (never a benchmark number): three warm training steps, then ONE step between
cudaProfilerStart / cudaProfilerStop, eager (no CUDA graph), side streams and PDL off so every kernel is measured alone.

    python bench/ncu_step.py --model transformer                      # plain run first (must exit 0)
    ncu --set full --clock-control none --profile-from-start off -o gpurun_out/step python bench/ncu_step.py
    ncu -i gpurun_out/step.ncu-rep --page raw --csv > gpurun_out/step_raw.csv   # read with bench/ncu_table.py
"""
import argparse
import os
import sys

os.environ.setdefault("B200ML_PDL", "0")
os.environ.setdefault("B200ML_WGRAD_STREAM", "0")
os.environ.setdefault("B200ML_OVERLAP_STREAMS", "0")
os.environ.setdefault("B200ML_OPT_OVERLAP", "0")

import torch  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.parallel.engine import DataParallelEngine  # noqa: E402


def build(name, dev):
    from b200ml.ops.loss import cross_entropy, linear_cross_entropy
    B = 32
    if name == "transformer":
        from b200ml.models.transformer import Transformer
        S = 200
        model = Transformer(512, 1024, 8, 0.1, 1, S, 19214, 10837, 19214).to(dev).train()
        eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)
        src = torch.randint(1, 10837, (B, S), device=dev)
        trg = torch.randint(1, 19214, (B, S), device=dev)
        src[:, 150:] = 0
        trg[:, 120:] = 0
        la = (torch.tril(torch.ones(S, S, device=dev)) == 0)[None, None]

        def step():
            eng.begin_step()
            h = model(src, trg, (src != 0)[:, None, None, :], la, la, return_hidden=True)
            linear_cross_entropy(h, model.linear.weight, model.linear.bias, trg, ignore_index=0).backward()
            eng.finish_step()
        return step
    from b200ml.models import LSTM, FashionMNISTModel, mnist_mlp
    if name.startswith("cnn"):
        ch, hw = (1, 28) if name == "cnn" else (3, 32)
        model = FashionMNISTModel(ch, 10, 10, image_size=hw).to(dev).train()
        eng = DataParallelEngine(model, optimizer="sgd", lr=0.01)
        x, y = torch.rand(B, ch, hw, hw, device=dev), torch.randint(0, 10, (B,), device=dev)
    elif name == "lstm":
        model = LSTM(95815, 32, 32, 4).to(dev).train()
        eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)
        x, y = torch.randint(1, 95815, (B, 129), device=dev), torch.randint(0, 4, (B,), device=dev)
    else:
        model = mnist_mlp().to(dev).train()
        eng = DataParallelEngine(model, optimizer="sgd", lr=0.03)
        x, y = torch.randn(B, 784, device=dev), torch.randint(0, 10, (B,), device=dev)
    hz = torch.zeros(2, B, 32, device=dev)

    def step():
        eng.begin_step()
        if name == "lstm":
            pred, _, _ = model(x, hz, hz)
            loss = cross_entropy(pred[:, -1, :], y)
        else:
            loss = cross_entropy(model(x), y)
        loss.backward()
        eng.finish_step()
    return step


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="transformer", choices=["transformer", "cnn", "cnn_cifar", "lstm", "mlp"])
    a = ap.parse_args()
    dev = torch.device("cuda")
    torch.manual_seed(0)
    step = build(a.model, dev)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    step()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("done")


if __name__ == "__main__":
    main()
