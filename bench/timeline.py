"""Diagnostic only (never a benchmark number): kernel timeline of the CUDA-graph-replayed training step bench.py times.

    python bench/timeline.py --out gpurun_out/timeline_n1.json                       # one GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29540 \
        bench/timeline.py --gpus 2 --out gpurun_out/timeline_n2.json                 # rank 0 writes

CUPTI activity records through torch.profiler: every kernel of a few graph replays with its stream, start and
duration.  The summary printed at the end answers "what is on the critical path": per-stream busy time, the span of
one step, idle gaps on the main stream, and which kernels run after the last backward kernel (the exposed tail of the
gradient all-reduce / optimizer).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import bench as B  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--allreduce", default="auto")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--out", default="gpurun_out/timeline.json")
    args = ap.parse_args()
    rank, world, local = B.dist_setup(args.gpus)
    ob = B.OwnBench(args, rank, world, local)
    for i in range(args.warmup):
        ob.step(i, False)
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(args.steps):
            ob.step(i, False)
        torch.cuda.synchronize()
    tmp = args.out + f".rank{rank}.trace.json"
    os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
    prof.export_chrome_trace(tmp)
    ev = json.load(open(tmp))["traceEvents"]
    os.remove(tmp)
    ks = [e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e]
    ks.sort(key=lambda e: e["ts"])
    t0 = ks[0]["ts"]
    rows = [{"name": e["name"][:90], "stream": e.get("args", {}).get("stream", e.get("tid")), "start": round(e["ts"] - t0, 2),
             "dur": round(e["dur"], 2)} for e in ks]
    if rank == 0:
        json.dump({"world": world, "steps": args.steps, "kernels": rows}, open(args.out, "w"))
        summarize(rows, args.steps)
    if world > 1:
        torch.distributed.barrier()
    ob.engine.close()
    if world > 1:
        torch.distributed.destroy_process_group()


def summarize(rows, steps):
    # split into steps at the counters_advance kernel that opens begin_step
    starts = [i for i, r in enumerate(rows) if "counters_advance" in r["name"]]
    # two advance kernels per step (optimizer step at the start, RNG step at the end): keep every other one
    opens = starts[0::2] if len(starts) >= 2 * steps else starts
    if len(opens) < 2:
        print("could not split steps", len(starts))
        return
    lo, hi = opens[-2], opens[-1]
    step = rows[lo:hi]
    s0 = step[0]["start"]
    end = max(r["start"] + r["dur"] for r in step)
    print(f"step span {end - s0:.1f} us, {len(step)} kernels (next step starts at +{rows[hi]['start'] - s0:.1f})")
    streams = {}
    for r in step:
        streams.setdefault(r["stream"], []).append(r)
    for sid, rs in sorted(streams.items(), key=lambda kv: -sum(r["dur"] for r in kv[1])):
        busy = sum(r["dur"] for r in rs)
        print(f"  stream {sid}: {len(rs):3d} kernels, busy {busy:8.1f} us, first +{rs[0]['start'] - s0:7.1f}, "
              f"last end +{rs[-1]['start'] + rs[-1]['dur'] - s0:7.1f}")
    print("timeline (start, dur, stream, name):")
    for r in step:
        print(f"  +{r['start'] - s0:8.1f} {r['dur']:7.1f}  s{r['stream']}  {r['name'][:70]}")


if __name__ == "__main__":
    main()
