import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests"))
import dist_worker as W
rank, world, dev = W.setup()
from b200ml.parallel.engine import DataParallelEngine
outs = {}
for m in ("fused", "nccl"):
    model = W.small_model(dev)
    eng = DataParallelEngine(model, optimizer="adam", lr=1e-3, allreduce=m, bucket_mb=0.05)
    W.train(model, eng, rank, world, dev, steps=3, global_batch=8, shard=True)
    eng.weights_hash()
    outs[m] = {k: v.detach().clone() for k, v in model.state_dict().items()}
    if rank == 0 and m == "fused":
        print("buckets", [(s, n, eng.comm.owned_slice(s, n)) for s, n, _, _ in eng.buckets][:12])
    eng.close()
if rank == 0:
    for k in outs["fused"]:
        e = (outs["fused"][k] - outs["nccl"][k]).abs()
        if e.max().item() > 1e-6:
            idx = (e > 1e-6).nonzero()
            print(f"{k}: max err {e.max().item():.3e}, {idx.shape[0]} of {e.numel()} elements differ, first {idx[0].tolist()} last {idx[-1].tolist()}")
    print("ALGO", os.environ.get("B200ML_AR_ALGO"), "STREAMS", os.environ.get("B200ML_WGRAD_STREAMS"), "done")
dist.barrier()
dist.destroy_process_group()
