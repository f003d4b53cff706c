"""b200ml — a Blackwell (B200, sm_100a) native take on the "PyTorch + Apache Spark" training stack.

Importable as ``b200ml`` (alias package at the repo root; this directory keeps the name the build brief asked for).
Layers:  ops (sm_100a kernels behind autograd) -> models -> parallel (fused all-reduce DP engine)
-> spark (SparkConf / SparkSession / TorchDistributor / MLlib-style estimator) -> entry scripts at the repo root.
"""
__version__ = "0.1.0"

from . import _native  # noqa: F401


def native_available() -> bool:
    return _native.available()
