"""Samples/s of the small reference models — CNN (FashionMNIST-shape 1x28x28 and CIFAR-shape
3x32x32), LSTM classifier (V = 95 815, T = 129), MLP (784-dim preset) — for this framework and for the UNMODIFIED
reference classes (ast-extracted from baseline/_ref, stock torch, DistributedDataParallel + NCCL when N > 1).

    python bench/bench_models.py --model cnn --impl b200ml
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/bench_models.py --model lstm --impl reference

Synthetic data, batch 32 per GPU, device-timed (CUDA events), max over ranks, one JSON line on rank 0.
"""
import argparse
import ast
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def extract(path, cls, extra=None):
    src = open(os.path.join(ROOT, "baseline", "_ref", path)).read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.ClassDef) and n.name == cls)
    ns = {"torch": torch, "nn": torch.nn}
    ns.update(extra or {})
    exec(compile(ast.Module([node], []), path, "exec"), ns)
    return ns[cls]


def bench_small_model(name, impl, rank, world, local, steps=50, warmup=5, no_graph=False):
    """One device-timed region of ``steps`` training steps of a small reference model; the process group (world > 1)
    must exist.  Returns the result dict (every rank; the time is the max over ranks)."""
    class A:
        pass
    a = A()
    a.model, a.impl, a.steps, a.warmup, a.no_graph = name, impl, steps, warmup, no_graph
    dev = torch.device("cuda", local)
    torch.manual_seed(0)
    B = 32
    g = torch.Generator().manual_seed(rank)
    if a.model.startswith("cnn"):
        ch, hw = (1, 28) if a.model == "cnn" else (3, 32)
        xs = torch.rand(8, B, ch, hw, hw, generator=g).to(dev)
        ys = torch.randint(0, 10, (8, B), generator=g).to(dev)
    elif a.model == "lstm":
        V, T = 95815, 129
        xs = torch.randint(1, V, (8, B, T), generator=g).to(dev)
        ys = torch.randint(0, 4, (8, B), generator=g).to(dev)
    else:
        xs = torch.randn(8, B, 784, generator=g).to(dev)
        ys = torch.randint(0, 10, (8, B), generator=g).to(dev)

    if a.impl == "reference":
        from torch import nn
        if a.model.startswith("cnn"):
            if a.model == "cnn_cifar":
                return {"impl": "reference", "model": a.model,
                        "unavailable": "the reference FashionMNISTModel hard-codes hidden*7*7 and fails on 3x32x32 inputs"}
            model = extract("pytorch_cnn.py", "FashionMNISTModel")(1, 10, 10).to(dev)
            opt = torch.optim.SGD(model.parameters(), lr=0.01)
        elif a.model == "lstm":
            model = extract("pytorch_lstm.py", "LSTM", {"vocab": {"0": 0}})(95815, 32, 32, 4).to(dev)
            opt = torch.optim.Adam(model.parameters(), lr=1e-3)
        else:
            class RefMLP(nn.Module):  # the reference class is fixed at 4-5-4-3; same structure at the 784-dim preset
                def __init__(self):
                    super().__init__()
                    self.layer_1, self.layer_2, self.layer_3 = nn.Linear(784, 512), nn.Linear(512, 256), nn.Linear(256, 10)
                    self.sigmoid = nn.Sigmoid()

                def forward(self, x):
                    return self.layer_3(self.sigmoid(self.layer_2(self.sigmoid(self.layer_1(x)))))
            model = RefMLP().to(dev)
            opt = torch.optim.SGD(model.parameters(), lr=0.03)
        net = model
        if world > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            net = DDP(model, device_ids=[local])
        loss_fn = nn.CrossEntropyLoss()
        model.train()

        def step(i):
            x, y = xs[i % 8], ys[i % 8]
            if a.model == "lstm":
                h = torch.zeros(2, B, 32, device=dev)
                pred, _, _ = net(x, h, h.clone())
                loss = loss_fn(pred[:, -1, :], y)
            else:
                loss = loss_fn(net(x), y)
            opt.zero_grad()
            loss.backward()
            opt.step()
    else:
        from b200ml.models import LSTM, FashionMNISTModel, mnist_mlp
        from b200ml.ops.loss import cross_entropy
        from b200ml.parallel.engine import DataParallelEngine
        if a.model.startswith("cnn"):
            model = FashionMNISTModel(ch, 10, 10, image_size=hw).to(dev)
            eng = DataParallelEngine(model, optimizer="sgd", lr=0.01)
        elif a.model == "lstm":
            model = LSTM(95815, 32, 32, 4).to(dev)
            eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)
        else:
            model = mnist_mlp().to(dev)
            eng = DataParallelEngine(model, optimizer="sgd", lr=0.03)
        model.train()
        x_s, y_s = xs[0].clone(), ys[0].clone()
        hz = torch.zeros(2, B, 32, device=dev)

        def body():
            eng.begin_step()
            if a.model == "lstm":
                pred, _, _ = model(x_s, hz, hz)
                loss = cross_entropy(pred[:, -1, :], y_s)
            else:
                loss = cross_entropy(model(x_s), y_s)
            loss.backward()
            eng.finish_step()

        for _ in range(3):
            body()
        graph = None if a.no_graph else eng.capture(body)

        def step(i):
            x_s.copy_(xs[i % 8])
            y_s.copy_(ys[i % 8])
            if graph is not None:
                graph.replay()
            else:
                body()

    for i in range(max(a.warmup, 3)):
        step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    if a.impl != "reference":
        eng.close()
    return {"impl": a.impl, "model": a.model, "n_gpus": world, "batch_per_gpu": B, "steps": a.steps,
            "ms_per_step": ms / a.steps, "samples_per_s": B * world * a.steps / (ms / 1e3)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="cnn", choices=["cnn", "cnn_cifar", "lstm", "mlp"])
    ap.add_argument("--impl", default="b200ml", choices=["b200ml", "reference"])
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--no-graph", action="store_true")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    res = bench_small_model(a.model, a.impl, rank, world, local, a.steps, a.warmup, a.no_graph)
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
