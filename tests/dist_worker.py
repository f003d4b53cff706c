"""Worker launched by tests (torchrun or mp.spawn): data-parallel engine checks on gloo/CPU or NCCL+fused/GPU.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/dist_worker.py <case>
"""
import os
import sys

os.environ.setdefault("B200ML_SPARSE_ROWS", "1")   # opt-in feature: the tests exercise it

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def setup():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    use_cuda = torch.cuda.is_available() and torch.cuda.device_count() >= world and os.environ.get("B200ML_TEST_CPU") != "1"
    if use_cuda:
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
        dev = torch.device("cuda", torch.cuda.current_device())
    else:
        dist.init_process_group("gloo")
        dev = torch.device("cpu")
    return rank, world, dev


def small_model(dev):
    from b200ml.models.transformer import Transformer
    torch.manual_seed(7)
    # target vocabulary 2000: the decoder table and the head are 1 MB each, so with 1 MB buckets the engine takes the
    # two-shot (sliced) path for them and the one-shot path for the small buckets
    m = Transformer(128, 256, 2, 0.0, 1, 32, 2000, 80, 2000).to(dev)
    m.eval()  # SentenceEmbedding keeps the reference's hard-coded Dropout(0.1); eval() makes runs comparable
    return m


def train(model, engine, rank, world, dev, steps, global_batch, shard):
    from b200ml.ops.loss import linear_cross_entropy
    S = 32
    g = torch.Generator().manual_seed(99)
    losses = []
    la = (torch.tril(torch.ones(S, S, device=dev)) == 0)[None, None]
    for _ in range(steps):
        src = torch.randint(1, 80, (global_batch, S), generator=g)
        trg = torch.randint(1, 96, (global_batch, S), generator=g)
        if shard:
            per = global_batch // world
            src, trg = src[rank * per:(rank + 1) * per], trg[rank * per:(rank + 1) * per]
        src, trg = src.to(dev), trg.to(dev)
        engine.begin_step()
        h = model(src, trg, (src != 0)[:, None, None, :], la, la, return_hidden=True)
        loss = linear_cross_entropy(h, model.linear.weight, model.linear.bias, trg, ignore_index=0)
        loss.backward()
        engine.finish_step()
        losses.append(float(loss.detach()))
    return losses


def case_dp_equals_big_batch(rank, world, dev, mode):
    """N ranks x batch B with gradient averaging == 1 rank x batch N*B (same data), and ranks stay bit-identical."""
    from b200ml.parallel.engine import DataParallelEngine
    model = small_model(dev)
    ref_state = {k: v.clone() for k, v in model.state_dict().items()}
    # SGD: the update is linear in the gradient, so summation-order noise stays at rounding level (Adam's first
    # steps are sign-like and amplify it)
    eng = DataParallelEngine(model, optimizer="sgd", lr=0.1, allreduce=mode, bucket_mb=0.05)
    assert len(eng.buckets) > 2
    if eng.comm is not None:
        assert any(eng.comm.owned_slice(s, n) is not None for s, n, _, _ in eng.buckets), "no two-shot bucket"
    train(model, eng, rank, world, dev, steps=4, global_batch=8, shard=True)
    eng.weights_hash()
    dp = torch.cat([v.detach().reshape(-1).clone() for v in model.state_dict().values()])
    eng.close()
    # single-rank oracle with the full batch (every rank computes it redundantly, no communication)
    model2 = small_model(dev)
    model2.load_state_dict(ref_state)
    eng2 = DataParallelEngine(model2, optimizer="sgd", lr=0.1, sync_gradients=False)
    train(model2, eng2, rank, world, dev, steps=4, global_batch=8, shard=False)
    ref = torch.cat([v.detach().reshape(-1) for v in model2.state_dict().values()])
    err = (dp - ref).abs().max().item()
    scale = ref.abs().max().item()
    tol = 3e-2 if dev.type == "cuda" else 1e-4   # bf16 activations on GPU: per-rank rounding differs from full batch
    assert err <= tol * scale, f"DP != big batch: err {err} scale {scale}"
    print(f"[rank {rank}] dp_equals_big_batch ok mode={eng.allreduce_mode} err={err:.3e}")


def case_fused_matches_nccl(rank, world, dev, mode):
    """The fused all-reduce + optimizer gives the same weights as NCCL all-reduce + the flat optimizer kernel.
    The two paths round differently in the last bit (NCCL adds in another order than our rank-order sum, and the
    scale / update arithmetic is fused differently); SGD keeps that at rounding level, while Adam turns the noise of
    gradients that are mathematically zero (the key bias of every attention block) into +-lr steps - there only the
    bulk of the elements is held to the tight tolerance.  (The exact Adam arithmetic of every transport is checked in
    allreduce_raw.)"""
    from b200ml.parallel.engine import DataParallelEngine
    for optimizer, lr in (("sgd", 0.1), ("adam", 1e-3)):
        outs = []
        for m in ("fused", "nccl"):
            model = small_model(dev)
            eng = DataParallelEngine(model, optimizer=optimizer, lr=lr, allreduce=m, bucket_mb=0.05)
            if m == "fused" and dev.type == "cuda" and os.environ.get("B200ML_AR_ALGO", "0") in ("0", "4", "5"):
                # the target table sits in a two-shot bucket: its all-reduce skips the rows a peer's batch did not touch
                # (the batches use < 100 of the 2000 target ids), and must still equal the dense NCCL reduction
                assert "sparse-rows" in eng.allreduce_mode and len(eng._sparse) >= 1, eng.allreduce_mode
            train(model, eng, rank, world, dev, steps=3, global_batch=8, shard=True)
            eng.weights_hash()
            # state_dict() pulls the fp32 slices this rank does not own (bf16-only push), and is layout independent
            outs.append(torch.cat([v.detach().reshape(-1).clone() for v in model.state_dict().values()]))
            w16 = eng.flat.flat_w16.float()
            assert (w16 - eng.flat.flat_p).abs().max().item() <= 0.01 * eng.flat.flat_p.abs().max().item()
            eng.close()
        err = (outs[0] - outs[1]).abs()
        if optimizer == "sgd":
            assert err.max().item() < 1e-5, f"sgd: fused vs nccl differ by {err.max().item()}"
        else:
            assert (err > 1e-5).float().mean().item() < 0.01 and err.max().item() <= 3.5 * 3 * lr, \
                f"adam: fused vs nccl differ by {err.max().item()}, {(err > 1e-5).float().mean().item():.4f} of the elements"
    print(f"[rank {rank}] fused_matches_nccl ok err={err.max().item():.3e}")


def case_allreduce_raw(rank, world, dev, mode):
    """kind-0 reduce of random buffers vs torch.distributed.all_reduce; SGD and Adam updates through every transport
    (1 one-shot, 2 SM-issued pull, 4 bulk-async TMA ring, 5 the same with per-thread weight stores), on whole buffers
    and on sub-ranges so small that some ranks own an EMPTY slice (the `lo > n` clamp) or only a 4-element granule."""
    from b200ml.parallel.comm import FusedAllReduce, SymmetricAllocator
    from b200ml.parallel.flat import FlatParams
    torch.manual_seed(rank)
    ps = [torch.randn(1000, 333, device=dev).requires_grad_(), torch.randn(4099, device=dev).requires_grad_()]
    alloc = SymmetricAllocator(dev)
    flat = FlatParams(ps, allocator=alloc)
    dist.broadcast(flat.flat_p, src=0)
    flat.refresh_working_copy()
    m = torch.zeros_like(flat.flat_p)
    v = torch.zeros_like(flat.flat_p)
    comm = FusedAllReduce(flat, m, v, None, allocator=alloc)
    step = torch.ones(1, dtype=torch.int32, device=dev)
    N = flat.numel
    ranges = [(0, N), (64, 24), (128, 8), (0, 4100), (64, 4096 + 8), (1024, 1024 * world + 8)]
    for algo in (1, 2, 4, 5):
        for start, n in ranges:
            if algo in (4, 5) and (start % 8 or n % 8):
                continue                      # the bulk ring needs 8-element granularity (auto falls back to 2)
            flat.flat_g.copy_(torch.randn(N, device=dev))
            expect = flat.flat_g.clone()
            dist.all_reduce(expect)
            expect /= world
            out = torch.zeros_like(flat.flat_g)
            comm.allreduce_update(start, n, 0, 0.0, (0.9, 0.999), 1e-8, 0.0, step, out=out, algo=1)
            torch.cuda.synchronize()
            assert (out[start:start + n] - expect[start:start + n]).abs().max().item() < 1e-5
            assert start == 0 or out[:start].abs().max().item() == 0
            # SGD
            p0 = flat.flat_p.clone()
            comm.allreduce_update(start, n, 1, 0.5, (0.9, 0.999), 1e-8, 0.0, step, algo=algo)
            torch.cuda.synchronize()
            dist.barrier()
            want = p0.clone()
            want[start:start + n] -= 0.5 * expect[start:start + n]
            assert (flat.flat_p - want).abs().max().item() < 1e-5, f"sgd algo {algo} range {(start, n)}"
            sl = slice(start, start + n)
            assert (flat.flat_w16[sl].float() - want[sl]).abs().max().item() < 0.05, f"w16 algo {algo} range {(start, n)}"
            # Adam (first step from zero moments): p -= lr * g / (|g| + eps); m / v live on the slice owners only
            p0 = flat.flat_p.clone()
            m.zero_()
            v.zero_()
            comm.allreduce_update(start, n, 2, 1e-2, (0.9, 0.999), 1e-8, 0.0, step, algo=algo)
            torch.cuda.synchronize()
            dist.barrier()
            # Adam divides by |g| + eps, so where the mean gradient is nearly zero the last-bit difference between NCCL's
            # summation order and ours would be amplified: use our own (rank-order) mean, checked against NCCL above
            g = out[sl]
            want = p0.clone()
            want[sl] -= 1e-2 * g / (g.abs() + 1e-8)
            assert (flat.flat_p - want).abs().max().item() < 1e-4, f"adam algo {algo} range {(start, n)}"
            own = comm.owned_slice(start, n, sliced=algo != 1)
            lo, hi = (0, n) if own is None else own
            assert hi <= lo or (m[start + lo:start + hi] - 0.1 * g[lo:hi]).abs().max().item() < 1e-6
            outside = torch.cat([m[:start + lo], m[start + hi:], m.new_zeros(1)])
            assert outside.abs().max().item() == 0, f"m written outside the owned slice, algo {algo} range {(start, n)}"
            dist.barrier()
    comm.close()
    print(f"[rank {rank}] allreduce_raw ok")


def case_allreduce_nvls(rank, world, dev, mode):
    """NVLS (multimem.ld_reduce / multimem.st) two-shot update vs torch.distributed.all_reduce; skipped when the box has
    no multicast support."""
    from b200ml.parallel.comm import FusedAllReduce, SymmMemAllocator, make_allocator
    from b200ml.parallel.flat import FlatParams
    alloc = make_allocator(dev)
    if not isinstance(alloc, SymmMemAllocator):
        print(f"[rank {rank}] allreduce_nvls skipped (no multicast)")
        return
    torch.manual_seed(rank)
    ps = [torch.randn(1024, 333, device=dev).requires_grad_(), torch.randn(4099, device=dev).requires_grad_()]
    flat = FlatParams(ps, allocator=alloc)
    dist.broadcast(flat.flat_p, src=0)
    flat.refresh_working_copy()
    comm = FusedAllReduce(flat, None, None, None, allocator=alloc)
    assert "nvls" in comm.transport, comm.transport
    step = torch.ones(1, dtype=torch.int32, device=dev)
    for push in (1, 0):
        flat.flat_g.copy_(torch.randn(flat.numel, device=dev))
        expect = flat.flat_g.clone()
        dist.all_reduce(expect)
        expect /= world
        p0 = flat.flat_p.clone()
        comm.push_master = push
        comm.allreduce_update(0, flat.numel, 1, 0.5, (0.9, 0.999), 1e-8, 0.0, step, algo=3)
        torch.cuda.synchronize()
        dist.barrier()
        want = p0 - 0.5 * expect
        assert (flat.flat_w16.float() - want).abs().max().item() < 0.05, "bf16 working copy"
        if not push:
            comm.pull_master(0, flat.numel)
            torch.cuda.synchronize()
        assert (flat.flat_p - want).abs().max().item() < 1e-4, f"push={push}"
        dist.barrier()
    comm.close()
    print(f"[rank {rank}] allreduce_nvls ok")


def case_checkpoint_adam(rank, world, dev, mode):
    """Adam state of a two-shot (sliced) bucket lives only on the slice owners: a checkpoint written from rank 0 must
    still hold ALL of it, and resuming from it must continue exactly like the uninterrupted run."""
    import tempfile
    from b200ml.models.transformer import Transformer
    from b200ml.parallel.engine import DataParallelEngine
    from b200ml.utils.checkpoint import load_checkpoint, save_checkpoint

    def build():
        torch.manual_seed(11)
        m = Transformer(128, 256, 2, 0.0, 1, 32, 2000, 80, 2000).to(dev)   # 2000 x 128 fp32 tables/head: 1 MB slots
        m.eval()
        return m

    def run(model, eng, steps, seed):
        from b200ml.ops.loss import linear_cross_entropy
        g = torch.Generator().manual_seed(seed + rank)
        la = (torch.tril(torch.ones(32, 32, device=dev)) == 0)[None, None]
        for _ in range(steps):
            src = torch.randint(1, 80, (4, 32), generator=g).to(dev)
            trg = torch.randint(1, 2000, (4, 32), generator=g).to(dev)
            eng.begin_step()
            h = model(src, trg, (src != 0)[:, None, None, :], la, la, return_hidden=True)
            linear_cross_entropy(h, model.linear.weight, model.linear.bias, trg, ignore_index=0).backward()
            eng.finish_step()

    ref = build()
    e_ref = DataParallelEngine(ref, optimizer="adam", lr=1e-3, allreduce=mode, bucket_mb=4.0)
    if e_ref.comm is not None:
        assert any(e_ref.comm.owned_slice(s, n) is not None for s, n, _, _ in e_ref.buckets), "no sliced bucket in test"
    run(ref, e_ref, 3, 1)
    run(ref, e_ref, 2, 2)
    want = torch.cat([v.detach().reshape(-1).clone() for v in ref.state_dict().values()])
    e_ref.close()

    a = build()
    e_a = DataParallelEngine(a, optimizer="adam", lr=1e-3, allreduce=mode, bucket_mb=4.0)
    run(a, e_a, 3, 1)
    path = os.path.join(tempfile.gettempdir(), f"b200ml_ck_{os.environ.get('MASTER_PORT', '0')}.pt")
    save_checkpoint(path, a, e_a)
    dist.barrier()
    blob = torch.load(path, weights_only=False)
    nz = sum(int((st["v"] != 0).sum()) for st in blob["optimizer"]["state"].values())
    tot = sum(st["v"].numel() for st in blob["optimizer"]["state"].values())
    e_a.close()

    b = build()
    with torch.no_grad():
        for p in b.parameters():
            p.add_(1.0)    # the checkpoint must overwrite this
    e_b = DataParallelEngine(b, optimizer="adam", lr=1e-3, allreduce=mode, bucket_mb=4.0)
    load_checkpoint(path, b, e_b, map_location="cpu")
    run(b, e_b, 2, 2)
    got = torch.cat([v.detach().reshape(-1).clone() for v in b.state_dict().values()])
    e_b.close()
    err = (got - want).abs().max().item()
    assert err < 1e-5, f"resume differs from the uninterrupted run by {err}"
    # e_b.allreduce_mode ends in ":rider" when the small fp32 parameters ride in a two-shot launch: their Adam moments
    # are then sliced over the ranks too, and the gather above covered them
    print(f"[rank {rank}] checkpoint_adam ok err={err:.3e} nonzero_v={nz}/{tot} mode={e_b.allreduce_mode}")


def case_compat_ddp(rank, world, dev, mode):
    """The reference's loop vocabulary (DDP(model) + Adam(model.parameters()) + zero_grad / backward / step) over several
    ranks trains exactly like the engine driven directly, and every rank ends with the same weights."""
    from b200ml.ops.loss import linear_cross_entropy
    from b200ml.optim import Adam
    from b200ml.parallel import DistributedDataParallel as DDP
    from b200ml.parallel.engine import DataParallelEngine

    ref = small_model(dev)
    e = DataParallelEngine(ref, optimizer="adam", lr=1e-3, allreduce=mode, bucket_mb=1.0)
    train(ref, e, rank, world, dev, 3, 8, True)
    want = torch.cat([p.detach().reshape(-1).clone() for p in ref.parameters()])
    e.close()

    model = small_model(dev)
    ddp = DDP(model, bucket_cap_mb=1.0, allreduce=mode)
    opt = Adam(model.parameters(), lr=1e-3)
    S = 32
    g = torch.Generator().manual_seed(99)
    la = (torch.tril(torch.ones(S, S, device=dev)) == 0)[None, None]
    per = 8 // world
    for _ in range(3):
        src = torch.randint(1, 80, (8, S), generator=g)[rank * per:(rank + 1) * per].to(dev)
        trg = torch.randint(1, 96, (8, S), generator=g)[rank * per:(rank + 1) * per].to(dev)
        h = ddp(src, trg, (src != 0)[:, None, None, :], la, la, return_hidden=True)
        loss = linear_cross_entropy(h, model.linear.weight, model.linear.bias, trg, ignore_index=0)
        opt.zero_grad()                 # between forward and backward, as in the reference
        loss.backward()
        opt.step()
    ddp.engine.weights_hash()           # raises if the ranks diverged
    got = torch.cat([p.detach().reshape(-1).clone() for p in model.parameters()])
    err = (got - want).abs().max().item()
    assert err == 0.0, f"wrapper API differs from the engine API by {err}"
    ddp.close()
    print(f"[rank {rank}] compat_ddp ok")


def case_layout_mismatch(rank, world, dev, mode):
    """A rank that built another model must be caught by the constructor, on every rank, before any peer access."""
    from b200ml.models.transformer import Transformer
    from b200ml.parallel.engine import DataParallelEngine
    vocab = 96 if rank == 0 else 104
    m = Transformer(128, 256, 2, 0.0, 1, 32, vocab, 80, vocab).to(dev)
    try:
        DataParallelEngine(m, optimizer="sgd", lr=0.1, allreduce=mode)
    except RuntimeError as e:
        assert "different parameter layouts" in str(e), str(e)
        print(f"[rank {rank}] layout_mismatch ok")
        return
    raise AssertionError("mismatched layouts were accepted")


def case_mlpc(rank, world, dev, mode):
    """MLlib-style MLPC over row shards: native loss / gradient kernels writing into the symmetric float64 vector, one
    one-shot float64 all-reduce; equals the full-data float64 restatement, and every rank ends with the same bits."""
    from b200ml.spark import ml
    layers = [20, 33, 7]
    torch.manual_seed(5)
    rows = 128 * world * 3                       # whole blocks per rank: the block weights of shards and full data agree
    X = torch.randn(rows, layers[0], dtype=torch.float64)
    Y = torch.randint(0, layers[-1], (rows,))
    per = rows // world
    Xs, Ys = X[rank * per:(rank + 1) * per].numpy(), Y[rank * per:(rank + 1) * per].numpy()
    out = ml._fit_local(Xs, Ys, layers, 128, 12, 1e-9, "l-bfgs", 0.03, 9, None, str(dev))
    w = torch.as_tensor(out["weights"])
    gathered = [torch.zeros_like(w) for _ in range(world)]
    wd = w.to(dev)
    gl = [torch.zeros_like(wd) for _ in range(world)]
    dist.all_gather(gl, wd)
    assert all(torch.equal(g, gl[0]) for g in gl), "ranks ended with different weights"
    # the objective the distributed run reports == the full-data objective at the returned weights
    row_w = torch.full((rows,), 1.0 / 128, dtype=torch.float64)
    loss, _ = ml.loss_and_grad(w, layers, X, Y, row_w)
    full = loss.item() / (rows / 128)
    assert abs(full - out["objectiveHistory"][-1]) <= 1e-10 * max(1.0, abs(full)), (full, out["objectiveHistory"][-1])
    assert out["objectiveHistory"][-1] < out["objectiveHistory"][0]
    print(f"[rank {rank}] mlpc ok evals={out['functionEvaluations']} obj={full:.6f}")


CASES = {"compat_ddp": case_compat_ddp, "mlpc": case_mlpc, "layout_mismatch": case_layout_mismatch, "checkpoint_adam": case_checkpoint_adam,"allreduce_nvls": case_allreduce_nvls, "dp_equals_big_batch": case_dp_equals_big_batch, "fused_matches_nccl": case_fused_matches_nccl,
         "allreduce_raw": case_allreduce_raw}

if __name__ == "__main__":
    rank, world, dev = setup()
    mode = os.environ.get("B200ML_TEST_MODE", "auto")
    for name in sys.argv[1:]:
        CASES[name](rank, world, dev, mode)
        dist.barrier()
    dist.destroy_process_group()
