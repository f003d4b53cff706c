"""Shared plumbing of the repo-root entry scripts: device / process-group set-up inside a worker, device-side timers,
small-run switches.  The reference has no flags at all; the scripts keep its module-level constants as
defaults and accept ``B200ML_*`` environment overrides so tests and quick runs do not have to edit files:

    B200ML_QUICK=1      tiny datasets / few steps (CI)
    B200ML_EPOCHS=n     override the epoch count
    B200ML_STEPS=n      cap the number of batches per epoch
    B200ML_DEVICE=cpu   force CPU
"""
from __future__ import annotations

import json
import os
import sys
import time

import torch
import torch.distributed as dist


def env_int(name: str, default: int) -> int:
    v = os.environ.get(name)
    return int(v) if v not in (None, "") else default


def quick() -> bool:
    return os.environ.get("B200ML_QUICK") == "1"


def pick_device() -> torch.device:
    if os.environ.get("B200ML_DEVICE", "") == "cpu" or not torch.cuda.is_available():
        return torch.device("cpu")
    local = int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count()
    torch.cuda.set_device(local)
    return torch.device("cuda", local)


def init_worker():
    """Inside a TorchDistributor/torchrun worker: create the process group that matches the device
    (NCCL for GPUs, gloo for CPU — the reference hard-codes 'gloo', distributed_cnn.py:152)."""
    dev = pick_device()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if dev.type == "cuda":
            dist.init_process_group("nccl", device_id=dev)
        else:
            dist.init_process_group("gloo")
    rank = dist.get_rank() if dist.is_initialized() else 0
    return dev, rank, world


def finish_worker():
    if dist.is_initialized():
        dist.destroy_process_group()


class DeviceTimer:
    """CUDA-event timer on the current stream (wall clock on CPU); never forces a sync until ``elapsed`` is read."""

    def __init__(self, device):
        self.cuda = torch.device(device).type == "cuda"
        self.t0 = None

    def start(self):
        if self.cuda:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        else:
            self.t0 = time.perf_counter()
        return self

    def elapsed(self) -> float:
        if self.cuda:
            self.e1.record()
            self.e1.synchronize()
            return self.e0.elapsed_time(self.e1) / 1e3
        return time.perf_counter() - self.t0


def max_over_ranks(seconds: float, device) -> float:
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([seconds], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return seconds


def log_json(rank: int, **fields) -> None:
    """Rank-0 structured log line."""
    if rank == 0:
        sys.stdout.write("B200ML_LOG " + json.dumps(fields) + "\n")
        sys.stdout.flush()
