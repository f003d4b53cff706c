"""All-reduce microbenchmark: the fused all-reduce + optimizer + bf16 re-cast kernels of
csrc/comm/allreduce.cu vs torch.distributed NCCL all_reduce (+ a separate optimizer kernel), at the gradient sizes of
the reference models: 31 KB (CNN), 12.3 MB (LSTM), 25 MiB (a DDP bucket), 116 MiB (Transformer-MT).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/bench_allreduce.py [--quick]

Device-timed (CUDA events), max over ranks.  busbw = 2(N-1)/N * bytes / t (NCCL's convention).  The roofline column
uses what actually crosses NVLink per direction in the fused two-shot update with bf16 weight push:
(N-1)/N * (4 + 2) bytes per parameter (fp32 gradients in, bf16 weights out), at the measured peer-copy rate.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.ops.optim import optim_step_flat  # noqa: E402
from b200ml.parallel.comm import FusedAllReduce, SymmetricAllocator  # noqa: E402
from b200ml.parallel.flat import FlatParams  # noqa: E402

SIZES = [("cnn_31KB", 7740), ("lstm_12.3MB", 3083108), ("bucket_25MiB", 25 * (1 << 20) // 4),
         ("transformer_116MiB", 30500110)]
PEER_COPY_GBS = 770.0   # measured cudaMemcpyPeer rate between two B200s of the box (round 1)


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
    quick = "--quick" in sys.argv
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    rows = []
    for name, n in SIZES:
        p = torch.randn(n, device=dev).requires_grad_()
        alloc = SymmetricAllocator(dev)
        flat = FlatParams([p], allocator=alloc)
        m = torch.zeros_like(flat.flat_p)
        v = torch.zeros_like(flat.flat_p)
        comm = FusedAllReduce(flat, m, v, None, allocator=alloc)
        step = torch.ones(1, dtype=torch.int32, device=dev)
        flat.flat_g.normal_()
        nbytes = flat.numel * 4
        big = nbytes >= (1 << 20)
        iters = 200 if not big else 30
        link_bytes = (world - 1) / world * 6 * flat.numel
        row = {"name": name, "bytes": nbytes, "world": world, "roofline_us_bf16push": link_bytes / PEER_COPY_GBS / 1e3}

        def record(key, ms):
            row[key + "_us"] = ms * 1e3
            row[key + "_busGBs"] = 2 * (world - 1) / world * nbytes / (ms * 1e-3) / 1e9

        variants = [("oneshot", 1, (0,)), ("pull", 2, (96, 148)), ("bulk", 4, (16, 24, 32, 48, 64, 96, 128, 148)),
                    ("bulkst", 5, (32, 64, 148))]
        if quick:
            variants = [("oneshot", 1, (0,)), ("pull", 2, (96,)), ("bulk", 4, (0, 32, 148))]
        for label, algo, blockset in variants:
            if algo == 1 and nbytes > (64 << 20):
                continue
            if algo != 1 and not big:
                blockset = (0,)
            for kind, kname in ((1, "sgd"), (2, "adam")):
                for blocks in blockset:
                    for push in ((1,) if algo == 1 else (0, 1) if (blocks in (0, 64, 96) and not quick) else (0,)):
                        comm.max_blocks = blocks
                        comm.push_master = push
                        ms = timed(lambda: comm.allreduce_update(0, flat.numel, kind, 1e-3, (0.9, 0.999), 1e-8, 0.0, step,
                                                                 algo=algo), iters, dev)
                        record(f"fused_{label}_{kname}" + (f"_b{blocks}" if blocks else "") + ("" if push else "_bf16push"), ms)
        comm.close()
        # NVLS variant (needs torch symmetric memory + multicast support)
        try:
            from b200ml.parallel.comm import SymmMemAllocator, make_allocator
            os.environ["B200ML_NVLS"] = "1"
            alloc2 = make_allocator(dev)
            if isinstance(alloc2, SymmMemAllocator) and big:
                p2 = torch.randn(n, device=dev).requires_grad_()
                flat2 = FlatParams([p2], allocator=alloc2)
                m2, v2 = torch.zeros_like(flat2.flat_p), torch.zeros_like(flat2.flat_p)
                comm2 = FusedAllReduce(flat2, m2, v2, None, allocator=alloc2)
                flat2.flat_g.normal_()
                for lab2, algo2, blockset in (("nvls", 3, (16, 32, 64, 148)), ("vmm_bulk", 4, (32, 148))):
                    for kind, kname in ((1, "sgd"), (2, "adam")):
                        for blocks in blockset:
                            comm2.max_blocks, comm2.push_master = blocks, 0
                            ms = timed(lambda: comm2.allreduce_update(0, flat2.numel, kind, 1e-3, (0.9, 0.999), 1e-8,
                                                                      0.0, step, algo=algo2), iters, dev)
                            record(f"fused_{lab2}_{kname}_b{blocks}_bf16push", ms)
                comm2.close()
        except Exception as e:  # noqa: BLE001
            row["nvls_error"] = str(e)[:200]
        g = flat.flat_g

        for kind, kname in ((1, "sgd"), (2, "adam")):
            def nccl_step():
                dist.all_reduce(g)
                optim_step_flat(flat.flat_p, flat.flat_g, m if kind == 2 else None, v if kind == 2 else None,
                                flat.flat_w16, kind, 1e-3, (0.9, 0.999), 1e-8, 0.0, step, 1.0 / world)
            record(f"nccl_plus_{kname}", timed(nccl_step, iters, dev))
        record("nccl_only", timed(lambda: dist.all_reduce(g), iters, dev))
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
