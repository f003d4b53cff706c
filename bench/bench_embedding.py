"""Embedding backward micro-benchmark (diagnostic): the sorted deterministic path at the translator's shapes.

    python bench/bench_embedding.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as B  # noqa: E402
from b200ml.ops.embedding import embedding  # noqa: E402
from b200ml.ops.optim import FlatOptimizer  # noqa: E402


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    dev = torch.device("cuda")
    for V in (19214, 10837):
        src, trg = B.make_batches(1, 32, 200, V, V, seed=1)[0]
        tok = trg.to(dev)
        for pad, p in ((None, 0.1), (0, 0.1), (None, 0.0)):
            w = torch.nn.Parameter(torch.randn(V, 512, device=dev))
            opt = FlatOptimizer([w], kind="sgd", lr=0.0)
            out = embedding(tok, w, dropout_p=p, site=5, padding_idx=pad)
            g = torch.randn_like(out)

            def fwd_bwd():
                o = embedding(tok, w, dropout_p=p, site=5, padding_idx=pad)
                o.backward(g)

            def fwd_only():
                with torch.no_grad():
                    embedding(tok, w, dropout_p=p, site=5, padding_idx=pad)

            t_all, t_fwd = timed(fwd_bwd), timed(fwd_only)
            print(f"V={V} padding_idx={pad} p={p}: fwd+index+bwd {t_all:.1f} us, fwd alone {t_fwd:.1f} us")


if __name__ == "__main__":
    main()
