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

dev = torch.device("cuda")
torch.manual_seed(0)
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
