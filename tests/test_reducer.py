"""Native gradient bucket manager (csrc/runtime/reducer.cpp) against its Python statement of the same contract."""
import random

import pytest


def _table(r):
    return [(s, n, tuple(ids), g) for s, n, ids, g in r.buckets]


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_native_reducer_matches_python(seed):
    from b200ml.parallel.reducer import NativeReducer, PyReducer
    rng = random.Random(seed)
    n = rng.randint(1, 40)
    sizes = [rng.choice([64, 128, 4096, 65536, 1 << 20]) for _ in range(n)]
    offsets = [0]
    for s in sizes:
        offsets.append(offsets[-1] + s)
    split = rng.randint(0, n)
    cap = rng.choice([1, 5000, 1 << 18, 1 << 30])
    nat, py = NativeReducer(offsets, [split, n], cap), PyReducer(offsets, [split, n], cap)
    assert _table(nat) == _table(py)
    # buckets tile the flat buffer, never straddle the group boundary
    covered = 0
    for s, m, ids, g in nat.buckets:
        assert s == covered and m == sum(sizes[i] for i in ids)
        assert all((i < split) == (g == 0) for i in ids)
        covered += m
    assert covered == offsets[-1]
    for p in range(n):
        assert nat.bucket_of(p) == py.bucket_of(p)
    for _ in range(2):   # two steps: the second checks begin_step() resets the state
        nat.begin_step(); py.begin_step()
        order = list(range(n))
        rng.shuffle(order)
        skipped = set(order[: n // 4])          # parameters unused in this forward
        fired = []
        for p in order:
            if p in skipped:
                continue
            a, b = nat.mark_ready(p), py.mark_ready(p)
            assert a == b
            assert nat.mark_ready(p) == -1 and py.mark_ready(p) == -1   # a second report never fires a bucket
            if a >= 0:
                fired.append(a)
        rest_n, rest_p = nat.flush(), py.flush()
        assert rest_n == rest_p
        assert sorted(fired + rest_n) == list(range(len(nat.buckets)))
        assert nat.flush() == []
    nat.close()


def test_engine_uses_native_reducer():
    import torch
    from b200ml.parallel.engine import DataParallelEngine
    from b200ml.parallel.reducer import NativeReducer
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4))
    eng = DataParallelEngine(model, optimizer="sgd", lr=0.1, bucket_mb=0.0001)
    assert isinstance(eng.reducer, NativeReducer)
    assert len(eng.buckets) >= 2
    eng.begin_step()
    loss = model(torch.randn(5, 8)).square().mean()
    loss.backward()
    eng.finish_step()
    eng.close()


def test_graphed_step_is_plain_call_on_cpu_and_close_is_idempotent():
    import pickle

    import torch
    from b200ml.parallel.engine import DataParallelEngine
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Sigmoid(), torch.nn.Linear(16, 3))
    eng = DataParallelEngine(model, optimizer="sgd", lr=0.1)
    calls = []

    def one_step(x, y):
        calls.append(tuple(x.shape))
        loss = torch.nn.functional.cross_entropy(model(x), y)
        eng.zero_grad()
        loss.backward()
        eng.step()
        return loss.detach()

    step = eng.graphed(one_step)
    assert not step.enabled                      # no CUDA device: every call is the eager function
    l0 = float(step(torch.randn(5, 8), torch.randint(0, 3, (5,))))
    for _ in range(20):
        l1 = float(step(torch.randn(5, 8), torch.randint(0, 3, (5,))))
    assert len(calls) == 21 and step.replays == 0 and l1 == l1 and l0 == l0
    eng.close()
    eng.close()                                  # a second close must be harmless
    pickle.loads(pickle.dumps(model))            # the module carries no engine state afterwards


def test_big_parameters_start_their_own_bucket():
    """A parameter of cap/4 or more never joins the parameters registered before it (the translator's output projection,
    ready at the very start of backward, must not wait for the decoder layer) but takes up to cap/2 of small followers
    along (an embedding table and the weights of the layer it feeds become ready together, at the end)."""
    from b200ml.parallel.reducer import NativeReducer, PyReducer
    M = 1 << 20
    sizes = [5 * M, M // 2, M // 4, 9 * M, M // 2, 9 * M, 64]      # emb, small, small, emb, small, head, bias
    offsets = [0]
    for s in sizes:
        offsets.append(offsets[-1] + s)
    cap = 6 * M + M // 2
    for R in (NativeReducer, PyReducer):
        r = R(offsets, [len(sizes)], cap)
        assert [ids for _, _, ids, _ in r.buckets] == [[0, 1, 2], [3, 4], [5, 6]], R.__name__
        r.close()
        # followers beyond cap/2 are not absorbed: they pack among themselves
        r = R([0, 9 * M, 10 * M, 11 * M, 12 * M, 13 * M, 22 * M], [6], cap)
        assert [ids for _, _, ids, _ in r.buckets] == [[0, 1, 2, 3], [4], [5]], R.__name__
        r.close()


def test_small_bucket_rides_with_the_next_two_shot_bucket(monkeypatch):
    """Scheduling of the riding bucket (DataParallelEngine._setup_rider): once ready it waits for the next two-shot
    bucket and is passed along as that launch's tail; when no two-shot bucket follows it in the step it is launched
    alone at the end, through the sliced (algo 4) path, so the owners of its optimizer state never change.  The
    communicator and the streams are stand-ins: only the engine's decisions are under test."""
    import contextlib

    import torch
    from b200ml.models.transformer import Transformer
    from b200ml.parallel import engine as engine_mod
    from b200ml.parallel.engine import DataParallelEngine

    eng = DataParallelEngine(Transformer(64, 128, 2, 0.0, 1, 16, 50, 40, 50), optimizer="adam", lr=1e-3, bucket_mb=0.05)
    assert len(eng.buckets) >= 3, eng.buckets
    calls = []

    class FakeComm:
        push_master = 1

        def allreduce_update(self, start, n, kind, lr, betas, eps, wd, step, out=None, algo=0, sparse=None, tail=None):
            calls.append((start, n, tail, algo))

    class FakeStream:
        cuda_stream = 0

        def wait_stream(self, other):
            pass

        def wait_event(self, ev):
            pass

    fake = FakeStream()
    monkeypatch.setattr(engine_mod.torch.cuda, "stream", lambda st: contextlib.nullcontext())
    monkeypatch.setattr(engine_mod.torch.cuda, "current_stream", lambda *a, **k: fake)
    eng.world, eng.comm, eng.comm_stream, eng._opt_streams, eng._local_overlap = 2, FakeComm(), fake, [], False
    rider, host, other = len(eng.buckets) - 1, 0, 1
    eng._rider, eng._hosts = rider, {host}
    r_start, r_n = eng.buckets[rider][0], eng.buckets[rider][1]

    # rider first, then a bucket that cannot carry it, then the host: ONE launch carries it
    eng._rider_waiting = False
    eng._bucket_ready(rider)
    assert calls == [] and eng._rider_waiting
    eng._bucket_ready(other)
    assert calls[-1][2] is None and eng._rider_waiting
    eng._bucket_ready(host)
    assert calls[-1][:2] == eng.buckets[host][:2] and calls[-1][2] == (r_start, r_n) and not eng._rider_waiting
    eng._flush_rider()
    assert len(calls) == 2

    # host first: the rider has nobody to ride with and goes alone at the end of the step, on the sliced path
    calls.clear()
    eng._bucket_ready(host)
    eng._bucket_ready(rider)
    assert len(calls) == 1 and calls[0][2] is None and eng._rider_waiting
    eng._flush_rider()
    assert calls[-1] == (r_start, r_n, None, 4) and not eng._rider_waiting

    # an incomplete host bucket (flushed at the end because a parameter never reported) does not carry it
    calls.clear()
    eng._bucket_ready(rider)
    eng._bucket_ready(host, complete=False)
    assert calls[-1][2] is None and eng._rider_waiting
    eng._flush_rider()
    assert calls[-1] == (r_start, r_n, None, 4)
    eng.comm = None
    eng.world = 1
    eng.close()


def _toy_batches(n, seed):
    import torch
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(16, 784, generator=g), torch.randint(0, 10, (16,), generator=g)) for _ in range(n)]


def test_reference_style_ddp_and_optimizer_api_matches_the_engine():
    """``DDP(model)`` + ``SGD / Adam(model.parameters(), lr)`` + ``zero_grad / backward / step`` (the reference's loop,
    distributed_cnn.py:156,138,175-183) must train exactly like driving the engine directly - with ``zero_grad()``
    before the forward pass or between forward and backward, and with the forward through the wrapper or the bare
    module (the reference forwards through the bare one)."""
    import torch
    from b200ml.models import mnist_mlp
    from b200ml.ops.loss import cross_entropy
    from b200ml.optim import SGD, Adam
    from b200ml.parallel import DistributedDataParallel as DDP
    from b200ml.parallel.engine import DataParallelEngine

    for kind, Opt, lr in (("sgd", SGD, 0.05), ("adam", Adam, 1e-3)):
        torch.manual_seed(0)
        ref = mnist_mlp()
        eng = DataParallelEngine(ref, optimizer=kind, lr=lr)
        for x, y in _toy_batches(5, 1):
            eng.zero_grad()
            cross_entropy(ref(x), y).backward()
            eng.step()
        want = torch.cat([p.detach().reshape(-1) for p in ref.parameters()])
        eng.close()
        for style in ("zero_first_wrapper_forward", "zero_between_bare_forward", "no_zero_wrapper_forward"):
            torch.manual_seed(0)
            model = mnist_mlp()
            ddp = DDP(model)
            opt = Opt(model.parameters(), lr=lr)          # on the BARE model's parameters, as the reference does
            assert opt.param_groups[0]["lr"] == lr and ddp.engine is not None
            for x, y in _toy_batches(5, 1):
                if style == "zero_first_wrapper_forward":
                    opt.zero_grad()
                    loss = cross_entropy(ddp(x), y)
                elif style == "zero_between_bare_forward":
                    loss = cross_entropy(model(x), y)
                    opt.zero_grad()
                else:
                    loss = cross_entropy(ddp(x), y)
                loss.backward()
                opt.step()
            got = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
            assert torch.equal(got, want), (kind, style, (got - want).abs().max().item())
            assert set(ddp.state_dict()) == {"module." + k for k in model.state_dict()}    # torch DDP's key prefix
            ddp.close()
    # without a wrapper the same optimizer classes drive the single-process flat optimizer
    torch.manual_seed(0)
    model = mnist_mlp()
    opt = SGD(model.parameters(), lr=0.05)
    for x, y in _toy_batches(5, 1):
        opt.zero_grad()
        cross_entropy(model(x), y).backward()
        opt.step()
    torch.manual_seed(0)
    ref = mnist_mlp()
    eng = DataParallelEngine(ref, optimizer="sgd", lr=0.05)
    for x, y in _toy_batches(5, 1):
        eng.zero_grad()
        cross_entropy(ref(x), y).backward()
        eng.step()
    a = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    b = torch.cat([p.detach().reshape(-1) for p in ref.parameters()])
    assert torch.allclose(a, b, atol=1e-6)
    import pytest as _pt
    with _pt.raises(ValueError):
        SGD(list(mnist_mlp().parameters()), lr=0.1, momentum=0.9)
    d2 = DDP(mnist_mlp())
    with _pt.raises(ValueError):
        Adam(list(d2.module.parameters())[:2], lr=1e-3)   # a subset of the wrapped module's parameters
