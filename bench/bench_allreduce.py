"""All-reduce microbenchmark: bus GB/s of the fused P2P kernels (one-shot / two-shot, with the
SGD update and bf16 re-cast folded in) vs torch.distributed NCCL all_reduce (+ separate optimizer kernel), at the
gradient sizes of the reference models: 31 KB (CNN), 12.3 MB (LSTM), 25 MiB (a DDP bucket), 116 MiB (Transformer-MT).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/bench_allreduce.py

Device-timed (CUDA events), max over ranks.  busbw = 2(N-1)/N * bytes / t.
"""
import json
import os
import sys

os.environ.setdefault("B200ML_AR_PUSH", "1")  # the benchmark measures every variant, including the opt-in ones

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.ops.optim import optim_step_flat  # noqa: E402
from b200ml.parallel.comm import FusedAllReduce, SymmetricAllocator  # noqa: E402
from b200ml.parallel.flat import FlatParams  # noqa: E402

SIZES = [("cnn_31KB", 7740), ("lstm_12.3MB", 3083108), ("bucket_25MiB", 25 * (1 << 20) // 4),
         ("transformer_116MiB", 30500110)]


def timed(fn, iters, dev):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    rows = []
    for name, n in SIZES:
        p = torch.randn(n, device=dev).requires_grad_()
        alloc = SymmetricAllocator(dev)
        flat = FlatParams([p], allocator=alloc)
        comm = FusedAllReduce(flat, None, None, None, allocator=alloc)
        step = torch.ones(1, dtype=torch.int32, device=dev)
        flat.flat_g.normal_()
        nbytes = flat.numel * 4
        iters = 200 if nbytes < (1 << 20) else 30
        row = {"name": name, "bytes": nbytes, "world": world}
        for label, algo in (("oneshot", 1), ("twoshot", 2), ("push", 4)):
            if algo == 1 and nbytes > (64 << 20):
                continue
            for blocks in ((0,) if nbytes < (1 << 20) else (64, 148, 296) if algo != 1 else (148,)):
                for push in ((1,) if algo == 1 else (1, 0)):
                    comm.max_blocks = blocks
                    comm.push_master = push
                    ms = timed(lambda: comm.allreduce_update(0, flat.numel, 1, 1e-3, (0.9, 0.999), 1e-8, 0.0, step,
                                                             algo=algo), iters, dev)
                    key = f"fused_{label}_sgd" + (f"_b{blocks}" if blocks else "") + ("" if push else "_bf16push")
                    row[key + "_us"] = ms * 1e3
                    row[key + "_busGBs"] = 2 * (world - 1) / world * nbytes / (ms * 1e-3) / 1e9
        comm.close()
        # NVLS variant (needs torch symmetric memory + multicast support)
        try:
            from b200ml.parallel.comm import SymmMemAllocator, make_allocator
            os.environ["B200ML_NVLS"] = "1"
            alloc2 = make_allocator(dev)
            if isinstance(alloc2, SymmMemAllocator) and nbytes >= (1 << 20):
                p2 = torch.randn(n, device=dev).requires_grad_()
                flat2 = FlatParams([p2], allocator=alloc2)
                comm2 = FusedAllReduce(flat2, None, None, None, allocator=alloc2)
                flat2.flat_g.normal_()
                for lab2, algo2 in (("nvls", 3), ("vmm_twoshot", 2), ("vmm_push", 4)):
                    if algo2 == 4 and not comm2.cs.stage[0]:
                        continue
                    for blocks in (148, 296):
                        for push in (1, 0):
                            comm2.max_blocks, comm2.push_master = blocks, push
                            ms = timed(lambda: comm2.allreduce_update(0, flat2.numel, 1, 1e-3, (0.9, 0.999), 1e-8, 0.0,
                                                                      step, algo=algo2), iters, dev)
                            key = f"fused_{lab2}_sgd_b{blocks}" + ("" if push else "_bf16push")
                            row[key + "_us"] = ms * 1e3
                            row[key + "_busGBs"] = 2 * (world - 1) / world * nbytes / (ms * 1e-3) / 1e9
                comm2.close()
        except Exception as e:  # noqa: BLE001
            row["nvls_error"] = str(e)[:200]
        g = flat.flat_g

        def nccl_step():
            dist.all_reduce(g)
            optim_step_flat(flat.flat_p, flat.flat_g, None, None, flat.flat_w16, 1, 1e-3, (0.9, 0.999), 1e-8, 0.0,
                            step, 1.0 / world)
        ms = timed(nccl_step, iters, dev)
        row["nccl_plus_sgd_us"] = ms * 1e3
        row["nccl_plus_sgd_busGBs"] = 2 * (world - 1) / world * nbytes / (ms * 1e-3) / 1e9
        ms = timed(lambda: dist.all_reduce(g), iters, dev)
        row["nccl_only_us"] = ms * 1e3
        row["nccl_only_busGBs"] = 2 * (world - 1) / world * nbytes / (ms * 1e-3) / 1e9
        if rank == 0:
            print(json.dumps(row), flush=True)
        rows.append(row)
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/bench_allreduce_{world}gpu.json", "w") as fh:
            json.dump(rows, fh, indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
