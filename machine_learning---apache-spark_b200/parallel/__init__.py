"""Data parallelism: flat parameter/gradient buckets, the fused P2P all-reduce(+optimizer) engine, launch helpers.

``DistributedDataParallel`` (+ ``b200ml.optim.SGD`` / ``Adam``) is the reference's ``DDP(model)`` /
``torch.optim.SGD(model.parameters(), lr)`` vocabulary on top of ``engine.DataParallelEngine`` (parallel/compat.py).
"""


def __getattr__(name):   # lazy: importing the package must not pull torch.distributed / the native library in
    if name == "DistributedDataParallel":
        from .compat import DistributedDataParallel

        return DistributedDataParallel
    if name == "DataParallelEngine":
        from .engine import DataParallelEngine

        return DataParallelEngine
    raise AttributeError(name)
