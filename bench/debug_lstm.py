import os, sys, torch
from torch.profiler import ProfilerActivity, profile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.models import LSTM
from b200ml.ops.loss import cross_entropy
from b200ml.parallel.engine import DataParallelEngine
dev = torch.device("cuda"); torch.manual_seed(0)
B, T, V = 32, 129, 95815
model = LSTM(V, 32, 32, 4).to(dev).train()
eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)
x = torch.randint(1, V, (B, T), device=dev); y = torch.randint(0, 4, (B,), device=dev)
hz = torch.zeros(2, B, 32, device=dev)
def body():
    eng.begin_step()
    pred, _, _ = model(x, hz, hz)
    loss = cross_entropy(pred[:, -1, :], y)
    loss.backward()
    eng.finish_step()
for _ in range(3): body()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3): body()
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / 3, e.count / 3) for e in prof.key_averages() if e.device_time_total > 0]
rows.sort(key=lambda r: -r[1])
print("sum kernel us/step", sum(r[1] for r in rows))
for k, t, c in rows[:25]: print(f"{t:9.1f} us x{c:5.1f} {k[:90]}")
# capture bisect
def try_capture(name, fn):
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g): fn()
        print("capture ok:", name)
    except Exception as e:
        print("capture FAILED:", name, str(e)[:120])
    torch.cuda.synchronize()
def fwd_only():
    with torch.no_grad(): model(x, hz, hz)
try_capture("forward(no grad)", fwd_only)
def fwd_bwd():
    pred, _, _ = model(x, hz, hz); cross_entropy(pred[:, -1, :], y).backward()
eng._in_step = False
try_capture("fwd+bwd (no engine)", fwd_bwd)
try_capture("full step", body)
