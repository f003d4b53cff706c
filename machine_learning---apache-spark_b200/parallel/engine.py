"""Data-parallel training engine — the role DDP(model) + optimizer.zero_grad()/step() play in the reference
(distributed_cnn.py:156,179-183; distributed_lstm.py:160,192-194; distributed_multilayer_perceptron.py:106,134-138),
rebuilt around flat buckets and a fused all-reduce + optimizer kernel.

Per step:
  begin_step()   bump the optimizer-step counter, clear the few gradient slots that accumulate
  forward/backward   wgrad kernels write straight into the flat fp32 gradient buffer; each parameter reports "ready"
  (as buckets complete)  world > 1: the bucket is reduced across ranks and the optimizer applied in ONE kernel
                         (parallel/comm.py: one-shot / two-shot P2P over NVLink), on a side stream, overlapping the
                         rest of backward;  "nccl" mode (the measured baseline) uses torch.distributed.all_reduce
  finish_step()  join the side stream (world == 1: one fused optimizer launch over the whole flat buffer), then bump
                 the device dropout-RNG step: forward and backward of one step always see the same draw, wherever
                 zero_grad() sits between them (the reference calls it after forward in distributed_lstm.py:192-194
                 and before forward in pytorch_machine_translator.py:172)

Rank 0's parameters are broadcast at construction (what DDP's constructor does); gradients are
averaged (sum / world), so N ranks with batch B reproduce one rank with batch N·B.

Weight traffic in the fused two-shot path: the GEMMs and the embedding gather only ever read the bf16 working copy of
nn.Linear / nn.Embedding weights (99.9 % of the Transformer's parameters), so for those buckets the slice owner pushes
just the bf16 values to its peers (2 B/param instead of 6) and keeps the fp32 master of its slice to itself; the other
ranks' fp32 views of such slices are refreshed on demand (``sync_master``: a one-sided pull over the same peer mappings,
run automatically before ``state_dict()``).  Biases, LayerNorm / conv / LSTM parameters are read in fp32 by their
kernels and stay fully replicated.
"""
from __future__ import annotations

import gc
import os

import torch
import torch.distributed as dist
from torch import nn

from .. import _native
from ..ops.optim import KINDS, optim_step_flat
from ..utils.rng import DeviceCounters
from .flat import FlatParams


class DataParallelEngine:
    def __init__(self, module, optimizer: str = "sgd", lr: float = 1e-2, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, allreduce: str = "auto", bucket_mb: float = 25.0,
                 process_group=None, sync_gradients: bool = True, push_master="auto"):
        self.module = module
        self.kind = KINDS[optimizer]
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.pg = process_group
        self.world = dist.get_world_size(self.pg) if dist.is_available() and dist.is_initialized() else 1
        # models may order the independent branches of their forward by it (models/transformer.py: which of two
        # concurrent branches is ISSUED first decides which gradient bucket backward completes first)
        module._b200_dp_world = self.world
        self.rank = dist.get_rank(self.pg) if self.world > 1 else 0
        self.sync_gradients = sync_gradients  # False emulates the reference as written
        params = [p for p in module.parameters() if p.requires_grad]
        self.device = params[0].device
        mode = "none"
        if self.world > 1 and self.sync_gradients:
            mode = allreduce
            if mode == "auto":
                mode = os.environ.get("B200ML_ALLREDUCE", "fused" if self.device.type == "cuda" else "nccl")
            if self.device.type != "cuda":
                mode = "gloo"
        sym_alloc = None
        if mode == "fused":
            from .comm import make_allocator

            sym_alloc = make_allocator(self.device, self.pg)

        # parameters whose kernels read only the bf16 working copy (nn.Linear / nn.Embedding weights) go first
        if push_master == "auto":
            push_master = os.environ.get("B200ML_PUSH_MASTER", "auto")
        self._bf16_push = mode == "fused" and str(push_master) in ("auto", "0", "False", "false")
        w16_only = set()
        if self._bf16_push:
            for m in module.modules():
                if isinstance(m, (nn.Linear, nn.Embedding)):
                    w16_only.add(id(m.weight))
        order = ([i for i, p in enumerate(params) if id(p) in w16_only] +
                 [i for i, p in enumerate(params) if id(p) not in w16_only])
        n_w16 = sum(1 for p in params if id(p) in w16_only)
        self.flat = FlatParams(params, allocator=sym_alloc, order=order)
        self.counters = DeviceCounters.get(self.device)
        self.step_tensor = torch.zeros(1, dtype=torch.int32, device=self.device)  # this engine's optimizer step
        self.m = torch.zeros_like(self.flat.flat_p) if self.kind == 2 else None
        self.v = torch.zeros_like(self.flat.flat_p) if self.kind == 2 else None
        self.launches_in_capture = 0
        self._graph = None
        self._graphed_steps: list = []
        self._sd_hook = None

        if self.world > 1:
            self._verify_layout()
            dist.broadcast(self.flat.flat_p, src=0, group=self.pg)
            self.flat.refresh_working_copy()

        # buckets: contiguous ranges of the flat buffer; gradients become ready roughly in reverse registration
        # order, so within each group the LAST bucket completes first.  (start, numel, [param indices], bf16_push)
        bucket_mb = float(os.environ.get("B200ML_BUCKET_MB", bucket_mb))   # experiment knob; the argument is the default
        cap = max(int(bucket_mb * (1 << 20) / 4), 1)
        n_params = len(self.flat.params)
        from .reducer import make_reducer

        # native bucket manager (csrc/runtime/reducer.cpp); group 0 = bf16-push weights, group 1 = everything else
        self.reducer = make_reducer(list(self.flat.offsets[:n_params]) + [self.flat.numel], [n_w16, n_params], cap)
        self.buckets = [(start, n, ids, group == 0 and n_w16 > 0) for start, n, ids, group in self.reducer.buckets]
        self._events = [dict() for _ in self.buckets]
        # side stream for the parameter-gradient kernels (ops/gemm.py GradSideStream)
        self.side = None
        if self.device.type == "cuda" and os.environ.get("B200ML_WGRAD_STREAM", "1") != "0":
            from ..ops.gemm import GradSideStream

            self.side = GradSideStream(self.device)
        for i, p in enumerate(self.flat.params):
            p._b200_ready = self._make_ready(i)
            p._b200_side = self.side

        self.allreduce_mode = mode
        self.comm = None
        self.comm_stream = None
        self._master_stale = False
        if self.world > 1 and self.sync_gradients:
            if mode == "fused":
                from .comm import FusedAllReduce

                self.comm = FusedAllReduce(self.flat, self.m, self.v, self.pg, allocator=sym_alloc)
                self.allreduce_mode = "fused:" + self.comm.transport + (":bf16push" if self._bf16_push else "")
                self._setup_sparse_rows()
                if self._bf16_push:
                    # model.state_dict() on any rank first pulls the fp32 slices it does not own
                    self._sd_hook = module.register_state_dict_pre_hook(lambda *a, **k: self.sync_master())
            if self.device.type == "cuda":
                self.comm_stream = torch.cuda.Stream(device=self.device, priority=-1)
        # single GPU: the per-bucket optimizer launches also run on a side stream, overlapping the rest of backward
        self._local_overlap = (self.world == 1 and self.device.type == "cuda" and self.sync_gradients and
                               os.environ.get("B200ML_OPT_OVERLAP", "1") != "0")
        # Single GPU: no cross-rank ordering to respect, so the per-bucket optimizer launches rotate over a few streams
        # and each waits only for ITS bucket's gradients (one stream would make a bucket that has been ready for
        # 100 us queue behind one whose last weight gradient is still being computed).  Multi-GPU: the fused kernels of
        # one communicator must run one at a time and in the same order on every rank -> one stream.
        self._opt_streams = []
        if self._local_overlap:
            n_opt = max(1, int(os.environ.get("B200ML_OPT_STREAMS", "3")))
            self._opt_streams = [torch.cuda.Stream(device=self.device) for _ in range(n_opt)]
            self.comm_stream = self._opt_streams[0]
        self._in_step = False
        self._streams_used: set = set()
        if not hasattr(self, "_sparse"):
            self._sparse = {}
        self._setup_rider()

    def _setup_rider(self) -> None:
        """The bucket of small fp32 parameters (biases, LayerNorm vectors: a few 10 KB) used to be a launch of its own -
        two cross-GPU barriers, 25-36 us, for next to no data, usually in the exposed tail of the step.  When the fused
        bulk-async transport is in use it RIDES in the launch of the first two-shot bucket that becomes ready after it
        (`b200ml_allreduce_update_tail`).  Its optimizer state is then sliced over the ranks like a two-shot bucket's -
        in every step, also when it has to be launched alone (no later two-shot bucket in that step) - so that the
        owners of the Adam moments never change.  `B200ML_AR_RIDER=0` keeps the separate one-shot launch."""
        self._rider = None            # bucket index of the riding bucket
        self._rider_waiting = False   # ready in this step, not launched yet
        self._hosts: set = set()      # buckets that may carry it
        if (self.comm is None or self.kind == 0 or os.environ.get("B200ML_AR_RIDER", "1") == "0"
                or not hasattr(self.comm, "rider_capable") or not self.comm.rider_capable()):
            return
        small = [b for b, (start, n, _, push) in enumerate(self.buckets)
                 if not push and self.comm.owned_slice(start, n) is None and start % 8 == 0]
        hosts = {b for b, (start, n, _, _) in enumerate(self.buckets)
                 if self.comm.owned_slice(start, n) is not None and start % 8 == 0 and n % 8 == 0}
        if len(small) == 1 and hosts:
            self._rider, self._hosts = small[0], hosts
            self.allreduce_mode += ":rider"

    def _owned_slice(self, b: int):
        """Slice of bucket b whose optimizer state this rank keeps (None: every rank keeps all of it)."""
        start, n, _, _ = self.buckets[b]
        return self.comm.owned_slice(start, n, sliced=True if b == self._rider else None)

    def _setup_sparse_rows(self) -> None:
        """Row-sparse tables (parameters an op flagged ``_b200_sparse_rows``: embedding tables, whose gradient rows are
        exactly zero for tokens the rank's batch does not contain).  For each two-shot bucket that holds one, every rank
        publishes a peer-mapped array of touched-row flags (written at forward time by the embedding's index kernel,
        ``param._b200_touched``), and the all-reduce fetches a peer's chunk of the table only if that peer touched it."""
        self._sparse = {}
        # Opt-in (B200ML_SPARSE_ROWS=1).  Measured on the translator at 8 GPUs it does not pay yet: the tail buckets are
        # bound by barrier skew and fixed costs rather than by bytes (1.608 ms with, 1.593 ms without; 2 GPUs: equal),
        # and the flag gathering at kernel start costs ~10 us per launch.  It is exact either way.
        if os.environ.get("B200ML_SPARSE_ROWS", "0") != "1" or not hasattr(self.comm, "make_sparse_rows"):
            return
        if self.comm.cs.mc_grads:          # NVLS transport reduces in the switch: always dense
            return
        for b, (start, n, ids, _) in enumerate(self.buckets):
            if self.comm.owned_slice(start, n) is None:
                continue
            for i in ids:
                p = self.flat.params[i]
                if getattr(p, "_b200_sparse_rows", False) and p.dim() == 2 and b not in self._sparse:
                    desc, flags = self.comm.make_sparse_rows(self.flat.offsets[i], p.shape[0], p.shape[1])
                    p._b200_touched = flags
                    self._sparse[b] = desc
        if self._sparse:
            self.allreduce_mode += ":sparse-rows"

    def _verify_layout(self) -> None:
        """Every rank must have built the same flat layout (what DDP's constructor verifies across ranks before it
        broadcasts, distributed_cnn.py:156 / SURVEY 2.4 C2).  The fused all-reduce dereferences peer pointers at
        identical offsets, so e.g. a vocabulary of another size on one rank would corrupt its peers' memory instead
        of failing: compare a digest of (offsets, shapes, dtypes, optimizer) first and raise on every rank."""
        import hashlib

        desc = repr((self.flat.numel, list(self.flat.offsets), [tuple(p.shape) for p in self.flat.params],
                     [str(p.dtype) for p in self.flat.params], self.kind))
        mine = hashlib.sha256(desc.encode()).hexdigest()
        seen = [None] * self.world
        dist.all_gather_object(seen, (mine, len(self.flat.params), self.flat.numel), group=self.pg)
        if any(s[0] != seen[0][0] for s in seen):
            detail = ", ".join(f"rank {r}: {s[1]} params / {s[2]} elements" for r, s in enumerate(seen))
            raise RuntimeError("DataParallelEngine: the ranks built different parameter layouts (" + detail +
                               "); every rank must construct the identical model")

    # ------------------------------------------------------------------------------------------------ bucket events
    def _make_ready(self, i):
        def ready():
            if not self._in_step:
                return
            b = self.reducer.bucket_of(i)
            if self.comm_stream is not None:
                # gradients of one bucket may be produced on several streams (backward runs each node on the stream of
                # its forward; wgrads run on the side stream): remember the latest point of each
                cur = torch.cuda.current_stream()
                ev = torch.cuda.Event()
                ev.record(cur)
                self._events[b][cur.cuda_stream] = ev
                if self.side is not None and self.side.active and self.side.used:
                    # deliver_grad queued this parameter's kernels on side.last just before reporting it ready
                    ev2 = torch.cuda.Event()
                    ev2.record(self.side.last)
                    self._events[b][self.side.last.cuda_stream] = ev2
            if self.reducer.mark_ready(i) == b:   # this report completed the bucket
                self._bucket_ready(b)
        return ready

    def _bucket_ready(self, b: int, complete: bool = True) -> None:
        if (self.world == 1 and not self._local_overlap) or not self.sync_gradients:
            return  # single fused optimizer launch in finish_step
        if self._rider is not None and complete and self.comm_stream is not None:
            if b == self._rider:
                self._rider_waiting = True      # launched with the next two-shot bucket (or alone in finish_step)
                return
        if self.comm_stream is not None:
            st = self._opt_streams[b % len(self._opt_streams)] if self._opt_streams else self.comm_stream
            self._streams_used.add(st)
            st.wait_stream(torch.cuda.current_stream())
            for ev in self._events[b].values():
                st.wait_event(ev)
            carry = self._rider_waiting and b in self._hosts and complete
            if carry:
                for ev in self._events[self._rider].values():
                    st.wait_event(ev)
                self._rider_waiting = False
            with torch.cuda.stream(st):
                self._reduce_and_update(b, complete, carry)
        else:
            self._reduce_and_update(b, complete)

    def _flush_rider(self) -> None:
        """End of the step: the riding bucket became ready after the last two-shot bucket - launch it alone."""
        if not self._rider_waiting:
            return
        self._rider_waiting = False
        rider = self._rider
        st = self.comm_stream
        self._streams_used.add(st)
        st.wait_stream(torch.cuda.current_stream())
        for ev in self._events[rider].values():
            st.wait_event(ev)
        with torch.cuda.stream(st):
            self._reduce_and_update(rider, True)

    def _reduce_and_update(self, b: int, complete: bool = True, carry: bool = False) -> None:
        start, n, _, bf16_push = self.buckets[b]
        if self.comm is not None:
            self.comm.push_master = 0 if (bf16_push and self._bf16_push) else 1
            tail = None
            if carry:
                tail = (self.buckets[self._rider][0], self.buckets[self._rider][1])
            # the riding bucket launched alone still takes the sliced (two-shot, bulk) path: see _setup_rider
            algo = 4 if b == self._rider else 0
            self.comm.allreduce_update(start, n, self.kind, self.lr, self.betas, self.eps, self.weight_decay,
                                       self.step_tensor, sparse=self._sparse.get(b) if complete else None, tail=tail,
                                       algo=algo)
            return
        if self.world == 1:
            optim_step_flat(self.flat.flat_p, self.flat.flat_g, self.m, self.v, self.flat.flat_w16, self.kind, self.lr,
                            self.betas, self.eps, self.weight_decay, self.step_tensor, 1.0, start, n)
            return
        g = self.flat.flat_g[start:start + n]
        dist.all_reduce(g, group=self.pg)
        optim_step_flat(self.flat.flat_p, self.flat.flat_g, self.m, self.v, self.flat.flat_w16, self.kind, self.lr,
                        self.betas, self.eps, self.weight_decay, self.step_tensor, 1.0 / self.world, start, n)

    # ------------------------------------------------------------------------------------------------ step API
    def begin_step(self) -> None:
        # only the optimizer step moves here; the dropout RNG step moves at the END of the step (finish_step), after
        # backward has regenerated its masks - zero_grad() may legally sit between forward and backward
        self.counters.advance(rng=False, opt_tensor=self.step_tensor)
        self.flat.zero_accumulating_slots()
        self.reducer.begin_step()
        for b in range(len(self.buckets)):
            self._events[b].clear()
        self._rider_waiting = False
        self._in_step = True
        if self.side is not None:
            self.side.active = True

    def finish_step(self) -> None:
        self._in_step = False
        if self.side is not None:
            self.side.active = False
            if self.comm_stream is None or not self.sync_gradients:
                self.side.join()
        if (self.world == 1 and not self._local_overlap) or not self.sync_gradients:
            optim_step_flat(self.flat.flat_p, self.flat.flat_g, self.m, self.v, self.flat.flat_w16, self.kind, self.lr,
                            self.betas, self.eps, self.weight_decay, self.step_tensor)
            self.counters.advance(rng=True)
            return
        # parameters that never reported (unused in this forward) still need their bucket processed
        for b in self.reducer.flush():
            self._bucket_ready(b, complete=False)
        self._flush_rider()
        if self.comm_stream is not None:
            # only the streams this step actually used: inside a graph capture, waiting on a stream that is not part
            # of the capture is an error
            for st in self._streams_used:
                torch.cuda.current_stream().wait_stream(st)
            self._streams_used.clear()
        if self.side is not None:
            self.side.join()
        # every backward kernel of this step is ordered before this point on the current stream (autograd joined its
        # streams, the side and comm streams were joined above): the next forward draws fresh dropout masks
        self.counters.advance(rng=True)
        self._master_stale = self._bf16_push and self.comm is not None

    # torch.optim-flavoured aliases so the entry scripts read like the reference loops
    def zero_grad(self, set_to_none: bool = True) -> None:
        self.begin_step()

    def step(self) -> None:
        self.finish_step()

    # ------------------------------------------------------------------------------------------------ CUDA graph
    def capture(self, step_fn):
        """Capture one full training step (forward, backward, all-reduce, optimizer) into a CUDA graph."""
        assert self.device.type == "cuda"
        if self.world > 1 and self.comm is None:
            return None  # NCCL baseline path stays eager
        # Autograd nodes left over from eager warm-up steps (reference cycles awaiting the collector) would keep their
        # AccumulateGrad nodes - created on the warm-up stream - alive, and backward under capture would then try to
        # synchronise the capture stream with that uncaptured stream.
        gc.collect()
        torch.cuda.synchronize()
        n0 = _native.launches()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            step_fn()
        self.launches_in_capture = _native.launches() - n0
        self._graph = g
        return g

    def graphed(self, step_fn, warmup: int = 2, max_graphs: int = 4):
        """Wrap a whole training step ``step_fn(*tensors) -> tensor(s)`` (zero_grad/forward/backward/step included) so
        that, once a given input signature has run ``warmup`` times eagerly, it is captured into a CUDA graph and
        replayed: inputs are copied into static buffers, the returned tensors are the graph's static outputs (valid
        until the next call).  Other signatures (e.g. a short last batch) and CPU runs stay eager.  ``B200ML_GRAPH=0``
        turns capturing off."""
        gs = GraphedStep(self, step_fn, warmup, max_graphs)
        self._graphed_steps.append(gs)
        return gs

    # ------------------------------------------------------------------------------------------------ utilities
    def sync_master(self) -> None:
        """Make every fp32 master parameter current on THIS rank (no-op unless the bf16-only push is active).
        One-sided: pulls the slices other ranks own through the peer mappings; safe to call on a single rank."""
        if self.comm is None or not self._bf16_push:
            return
        if not self._master_stale and self._graph is None:
            return
        torch.cuda.current_stream().synchronize()
        for b, (start, n, _, bf16_push) in enumerate(self.buckets):
            if bf16_push:
                self.comm.pull_master(start, n)
        torch.cuda.current_stream().synchronize()
        self._master_stale = False

    def weights_hash(self) -> float:
        """Cheap rank-divergence check: every rank must hold bit-identical master weights."""
        self.sync_master()
        h = self.flat.flat_p.double().sum()
        if self.world > 1:
            lo, hi = h.clone(), h.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.pg)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.pg)
            if (hi - lo).abs().item() != 0.0:
                raise RuntimeError(f"rank divergence: weight checksum differs across ranks ({lo.item()} vs {hi.item()})")
        return float(h.item())

    def gather_moments(self):
        """Full Adam moments (m, v), identical on every rank.  In the fused two-shot all-reduce only the owner of a
        slice runs the optimizer for it, so a rank's own ``m`` / ``v`` are current only inside the slices it owns; the
        other ranges are stale (zero if never loaded).  Collective: every rank must call it.  Each rank contributes
        its owned slices, the rest comes from a sum over ranks of buffers that are zero outside the owned ranges."""
        if self.m is None:
            return None, None
        if self.comm is None or self.world == 1:
            return self.m, self.v
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        out = []
        for t in (self.m, self.v):
            full = t.clone()
            for b, (start, n, _, _) in enumerate(self.buckets):
                own = self._owned_slice(b)
                if own is None:
                    continue        # one-shot bucket: every rank ran the identical update
                lo, hi = own
                n4 = (n + 3) // 4 * 4
                full[start:start + lo].zero_()
                full[start + hi:start + n4].zero_()
                dist.all_reduce(full[start:start + n4], group=self.pg)
            out.append(full)
        return out[0], out[1]

    def state_dict(self):
        """Collective when world > 1 (see gather_moments)."""
        m, v = self.gather_moments()
        return {"step": int(self.step_tensor.item()), "m": m, "v": v, "kind": self.kind}

    def load_state_dict(self, sd) -> None:
        self.step_tensor.fill_(sd["step"])
        if self.m is not None and sd.get("m") is not None:
            self.m.copy_(sd["m"])
            self.v.copy_(sd["v"])

    def close(self) -> None:
        # drop everything that is not picklable / must not outlive the engine: a script may return the module afterwards
        # (TorchDistributor ships rank 0's return value to the driver); graphs go before the peer mappings they use
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        for gs in self._graphed_steps:
            gs._graphs.clear()
            gs.enabled = False
        self._graphed_steps = []
        self._graph = None
        if self.comm is not None:
            self.sync_master()
            self.comm.close()
            self.comm = None
        if self._sd_hook is not None:
            self._sd_hook.remove()
            self._sd_hook = None
        self.reducer.close()
        for p in self.flat.params:
            if getattr(p, "_b200_touched", None) is not None:
                p._b200_touched = None
            p._b200_ready = None
            p._b200_side = None
        if getattr(self.module, "_b200_dp_world", None) is not None:
            self.module._b200_dp_world = 1


class GraphedStep:
    """See ``DataParallelEngine.graphed``.  The reference's loops call zero_grad / backward / step eagerly
    (pytorch_machine_translator.py:172-196); here the same loop body is replayed as one CUDA graph per input shape."""

    def __init__(self, engine: DataParallelEngine, step_fn, warmup: int = 2, max_graphs: int = 4):
        self.engine, self.fn, self.warmup, self.max_graphs = engine, step_fn, warmup, max_graphs
        self.enabled = engine.device.type == "cuda" and os.environ.get("B200ML_GRAPH", "1") != "0"
        self._seen: dict = {}
        self._graphs: dict = {}
        self.replays = 0

    def __call__(self, *inputs):
        if not self.enabled:
            return self.fn(*inputs)
        key = tuple((tuple(t.shape), t.dtype) for t in inputs)
        entry = self._graphs.get(key)
        if entry is None:
            n = self._seen.get(key, 0) + 1
            self._seen[key] = n
            if n <= self.warmup or len(self._graphs) >= self.max_graphs:
                return self.fn(*inputs)      # eager: first uses allocate workspaces / build caches
            static_in = [t.clone() for t in inputs]
            holder = []
            graph = self.engine.capture(lambda: holder.append(self.fn(*static_in)))
            if graph is None:                # e.g. the NCCL baseline path stays eager
                self.enabled = False
                return self.fn(*inputs)
            entry = (graph, static_in, holder[0])
            self._graphs[key] = entry
        graph, static_in, out = entry
        for s, t in zip(static_in, inputs):
            s.copy_(t, non_blocking=True)
        graph.replay()
        self.replays += 1
        return out
