"""MLlib-style MultilayerPerceptronClassifier / MulticlassClassificationEvaluator
(reference: mllib_multilayer_perceptron_classifier.py:1-2,32-48; SURVEY §3.4 restates the Spark 3.5 internals).

Math kept from MLlib: feed-forward topology Affine→Sigmoid→…→Affine→Softmax+cross-entropy; ONE flat float64 weight
vector laid out per layer as [W (out×in, column-major) ‖ b (out)], initialised U(−2.4, 2.4)/√in; rows stacked into blocks
of ``blockSize`` and the objective averaged per block, then over blocks; ``solver='l-bfgs'`` = L-BFGS with 10
corrections and a strong-Wolfe line search (``stepSize`` is ignored, as in MLlib — SURVEY §9 Q18), ``solver='gd'`` =
full-batch gradient descent with step ``stepSize/√iter``.

What changes on B200: every function evaluation is a forward/backward over the rank's shard of rows on its GPU followed
by ONE all-reduce of the vector [gradient ‖ loss ‖ #blocks] — the fused one-shot NVLink kernel of parallel/comm.py on
GPUs (gloo on CPU ranks) — instead of a Spark job (broadcast weights + treeAggregate).  The L-BFGS state is replicated:
every rank applies the identical update to identical data, so no weight broadcast exists at all.
"""
from __future__ import annotations

import math
import os
import pickle
import time

import numpy as np
import torch
import torch.distributed as dist

from .dataframe import DataFrame
from .linalg import DenseVector


# ---------------------------------------------------------------------------------------------------------------
# topology helpers
# ---------------------------------------------------------------------------------------------------------------
def num_weights(layers) -> int:
    return sum(layers[i] * layers[i + 1] + layers[i + 1] for i in range(len(layers) - 1))


def init_weights(layers, seed, dtype=torch.float64) -> torch.Tensor:
    g = torch.Generator().manual_seed(int(seed) & 0x7FFFFFFF)
    parts = []
    for i in range(len(layers) - 1):
        n_in, n_out = layers[i], layers[i + 1]
        n = n_in * n_out + n_out
        parts.append((torch.rand(n, generator=g, dtype=torch.float64) * 4.8 - 2.4) / math.sqrt(n_in))
    return torch.cat(parts).to(dtype)


def unpack(weights: torch.Tensor, layers):
    """Views (W [out,in], b [out]) per layer out of the flat MLlib-layout vector."""
    out, off = [], 0
    for i in range(len(layers) - 1):
        n_in, n_out = layers[i], layers[i + 1]
        W = weights[off:off + n_in * n_out].view(n_in, n_out).t()   # column-major (out×in) == row-major (in×out)
        off += n_in * n_out
        b = weights[off:off + n_out]
        off += n_out
        out.append((W, b))
    return out


def forward(weights, layers, X):
    """Returns (activations list, softmax probabilities)."""
    acts = [X]
    h = X
    params = unpack(weights, layers)
    for li, (W, b) in enumerate(params):
        z = h @ W.t() + b
        if li < len(params) - 1:
            h = torch.sigmoid(z)
            acts.append(h)
        else:
            return acts, torch.softmax(z, dim=1), z
    raise AssertionError


def loss_and_grad(weights, layers, X, Y, row_w):
    """Σ_rows row_w · CE(row) and its gradient w.r.t. the flat vector (manual backward, MLlib layout).
    ``row_w`` = 1/|block(row)| reproduces MLlib's per-block averaging."""
    acts, P, _ = forward(weights, layers, X)
    params = unpack(weights, layers)
    logp = torch.log(P.clamp_min(1e-300))
    loss = -(row_w * logp.gather(1, Y.unsqueeze(1)).squeeze(1)).sum()
    delta = P.clone()
    delta[torch.arange(X.shape[0], device=X.device), Y] -= 1.0
    delta = delta * row_w.unsqueeze(1)
    grads = []
    for li in range(len(params) - 1, -1, -1):
        W, _ = params[li]
        gW = delta.t() @ acts[li]                 # [out, in]
        gb = delta.sum(0)
        grads.append((gW, gb))
        if li > 0:
            a = acts[li]
            delta = (delta @ W) * a * (1 - a)
    grads.reverse()
    flat = torch.cat([torch.cat([gW.t().reshape(-1), gb]) for gW, gb in grads])
    return loss, flat


# ---------------------------------------------------------------------------------------------------------------
# L-BFGS (two-loop recursion + strong-Wolfe line search), replicated on every rank
# ---------------------------------------------------------------------------------------------------------------
def _native_direction(g, S, Yv):
    """K20 on the GPU: the whole two-loop recursion as one cooperative kernel (csrc/kernels/lbfgs.cu); float64 and
    bit-identical across ranks (fixed-order reductions).  Returns None when the problem exceeds the kernel's envelope."""
    import ctypes as C

    from .. import _native, _native_ext
    from ..utils import workspace

    vp, i, ll, sz = C.c_void_p, C.c_int, C.c_longlong, C.c_size_t
    _native_ext.register("b200ml_lbfgs_direction", [vp, vp, vp, i, ll, ll, vp, vp, sz, vp])
    _native_ext.register("b200ml_lbfgs_scratch_bytes", [], size_t_return=True)
    lib = _native.load()
    n = g.numel()
    props = torch.cuda.get_device_properties(g.device)
    if n > props.multi_processor_count * 256 * 16:
        return None
    g = g.contiguous()
    k = len(S)
    Sb = torch.stack(S).contiguous() if k else g
    Yb = torch.stack(Yv).contiguous() if k else g
    d = torch.empty_like(g)
    nb = int(lib.b200ml_lbfgs_scratch_bytes())
    ws = workspace.get(f"lbfgs@{workspace.lane()}", nb, g.device)
    _native.check(lib.b200ml_lbfgs_direction(g.data_ptr(), Sb.data_ptr(), Yb.data_ptr(), k, n, n, d.data_ptr(),
                                             ws.data_ptr(), nb, _native.stream_ptr()), "lbfgs_direction")
    return d


class LBFGS:
    def __init__(self, fn, m: int = 10, max_iter: int = 100, tol: float = 1e-6):
        self.fn, self.m, self.max_iter, self.tol = fn, m, max_iter, tol
        self.history: list[float] = []
        self.evals = 0

    def _f(self, w):
        self.evals += 1
        return self.fn(w)

    def two_loop(self, g, S, Yv):
        if g.is_cuda and g.dtype == torch.float64 and len(S) <= 32:
            d = _native_direction(g, S, Yv)
            if d is not None:
                return d
        q = g.clone()
        alphas = []
        for s, y in zip(reversed(S), reversed(Yv)):
            rho = 1.0 / torch.dot(y, s)
            a = rho * torch.dot(s, q)
            alphas.append((a, rho))
            q -= a * y
        if S:
            q *= torch.dot(S[-1], Yv[-1]) / torch.dot(Yv[-1], Yv[-1])
        for (a, rho), s, y in zip(reversed(alphas), S, Yv):
            b = rho * torch.dot(y, q)
            q += (a - b) * s
        return -q

    def _line_search(self, w, f0, g0, d, t0):
        """Strong Wolfe (c1=1e-4, c2=0.9): bracketing + zoom with cubic/bisection fallback."""
        c1, c2 = 1e-4, 0.9
        dg0 = float(torch.dot(g0, d))
        if dg0 >= 0:
            return 0.0, f0, g0

        def phi(t):
            f, g = self._f(w + t * d)
            return float(f), g, float(torch.dot(g, d))

        def zoom(lo, flo, dlo, hi, fhi):
            f = g = None
            t = lo
            for _ in range(30):
                t = 0.5 * (lo + hi)
                denom = 2.0 * (fhi - flo - dlo * (hi - lo))
                if denom != 0:  # quadratic interpolation, safeguarded
                    tq = lo - dlo * (hi - lo) ** 2 / denom
                    if min(lo, hi) + 0.1 * abs(hi - lo) < tq < max(lo, hi) - 0.1 * abs(hi - lo):
                        t = tq
                f, g, dg = phi(t)
                if f > f0 + c1 * t * dg0 or f >= flo:
                    hi, fhi = t, f
                else:
                    if abs(dg) <= -c2 * dg0:
                        return t, f, g
                    if dg * (hi - lo) >= 0:
                        hi, fhi = lo, flo
                    lo, flo, dlo = t, f, dg
                if abs(hi - lo) < 1e-16:
                    break
            return t, f, g

        t_prev, f_prev, d_prev = 0.0, f0, dg0
        t = t0
        f = g = None
        for i in range(25):
            f, g, dg = phi(t)
            if f > f0 + c1 * t * dg0 or (i > 0 and f >= f_prev):
                return zoom(t_prev, f_prev, d_prev, t, f)
            if abs(dg) <= -c2 * dg0:
                return t, f, g
            if dg >= 0:
                return zoom(t, f, dg, t_prev, f_prev)
            t_prev, f_prev, d_prev = t, f, dg
            t *= 2.0
        return t, f, g

    def minimize(self, w0):
        w = w0.clone()
        f, g = self._f(w)
        f = float(f)
        self.history = [f]
        S, Yv = [], []
        for it in range(self.max_iter):
            gnorm = float(g.norm())
            if gnorm <= self.tol * max(1.0, abs(f)):
                break
            d = self.two_loop(g, S, Yv)
            t0 = 1.0 if it > 0 else min(1.0, 1.0 / max(gnorm, 1e-12))
            t, f_new, g_new = self._line_search(w, f, g, d, t0)
            if t == 0.0 or g_new is None:
                break
            s = t * d
            y = g_new - g
            w = w + s
            if float(torch.dot(s, y)) > 1e-12 * float(y.norm()) * float(s.norm()):
                S.append(s)
                Yv.append(y)
                if len(S) > self.m:
                    S.pop(0)
                    Yv.pop(0)
            g = g_new
            self.history.append(f_new)
            # Breeze: function values converged (relative improvement over the recent window)
            if len(self.history) > 10:
                window = self.history[-11:]
                if abs(window[0] - window[-1]) <= self.tol * max(abs(window[0]), 1e-12):
                    f = f_new
                    break
            if abs(f - f_new) <= 1e-15 * max(1.0, abs(f)):
                f = f_new
                break
            f = f_new
        return w, it + 1 if self.history else 0


# ---------------------------------------------------------------------------------------------------------------
# distributed cost function
# ---------------------------------------------------------------------------------------------------------------
def _native_loss_grad(weights, layers, X, Y, row_w, out_vec):
    """K19 on the GPU (csrc/kernels/mlpc.cu): writes [gradient ‖ weighted loss] (num_weights + 1 float64) into
    ``out_vec`` with the float64 GEMM / softmax kernels of this library - no torch / cuBLAS op in the evaluation."""
    import ctypes as C

    from .. import _native, _native_ext
    from ..utils import workspace

    vp, i, ll, sz = C.c_void_p, C.c_int, C.c_longlong, C.c_size_t
    _native_ext.register("b200ml_mlpc_loss_grad", [vp, vp, i, vp, ll, vp, vp, i, vp, vp, sz, vp])
    _native_ext.register("b200ml_mlpc_scratch_bytes", [vp, i, i], size_t_return=True)
    lib = _native.load()
    la = (C.c_int * len(layers))(*[int(v) for v in layers])
    rows = X.shape[0]
    nb = int(lib.b200ml_mlpc_scratch_bytes(la, len(layers), rows))
    ws = workspace.get(f"mlpc@{workspace.lane()}", nb, X.device)
    _native.check(lib.b200ml_mlpc_loss_grad(weights.data_ptr(), la, len(layers), X.data_ptr(), X.stride(0),
                                            Y.data_ptr(), row_w.data_ptr(), rows, out_vec.data_ptr(), ws.data_ptr(),
                                            ws.numel(), _native.stream_ptr()), "mlpc_loss_grad")


class _Reducer:
    """Owner of the [grad ‖ loss ‖ blocks] vector and its sum over ranks.  GPUs of one node: the vector lives in a
    symmetric (peer-mapped) float64 buffer that the loss / gradient kernels write directly, and ONE one-shot kernel
    adds it over the ranks in rank order (parallel/comm.py F64AllReduce) - native float64, no hi/lo split, no staging
    copies.  CPU ranks (gloo) or ``B200ML_ALLREDUCE=nccl``: torch.distributed.all_reduce."""

    def __init__(self, n: int, device, dtype):
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.fused = None
        self.n = n
        if self.world > 1 and device.type == "cuda" and os.environ.get("B200ML_ALLREDUCE", "fused") == "fused":
            from ..parallel.comm import F64AllReduce

            self.fused = F64AllReduce(n, device)
            self.vec = self.fused.buf[:n]
        else:
            self.vec = torch.zeros(n, dtype=dtype, device=device)

    def __call__(self) -> torch.Tensor:
        """Sum of ``self.vec`` over the ranks (``self.vec`` itself when there is one rank)."""
        if self.world == 1:
            return self.vec
        if self.fused is not None:
            return self.fused.reduce()[:self.n]
        out = self.vec.clone()
        dist.all_reduce(out)
        return out

    def close(self):
        if self.fused is not None:
            self.fused.close()


def _block_weights(n: int, block_size: int, dev):
    """MLlib DataStacker: consecutive rows form blocks of <= blockSize; a row weighs 1 / (rows of its block)."""
    sizes = torch.full((n,), float(block_size), dtype=torch.float64, device=dev)
    rem = n % block_size
    if rem:
        sizes[n - rem:] = float(rem)
    return 1.0 / sizes


def _chunked_fn(Xd, Yd, layers, block_size, chunk_rows, chunks_local, world, native):
    """Loss / gradient evaluation whose floating-point result does not depend on the number of workers
    (``reproducible=True``).  The data set is cut into CANONICAL_CHUNKS row chunks (whole MLlib blocks each); a worker
    owns a run of consecutive chunks, evaluates [gradient ‖ loss ‖ #blocks] per chunk with the same code on the same
    rows whatever the worker count, the per-chunk vectors of all workers are all-gathered in chunk order and added
    pairwise in a fixed tree.  Every addition therefore has the same operands in the same order for 1, 2, 4 or 8
    workers: objective history and weights are bit-identical (tests/test_spark_and_data.py)."""
    dev, n, nw = Xd.device, Xd.shape[0], num_weights(layers)
    vecs = torch.zeros(chunks_local, nw + 2, dtype=torch.float64, device=dev)
    parts = []
    for c in range(chunks_local):
        a, b = min(c * chunk_rows, n), min((c + 1) * chunk_rows, n)
        if b > a:
            # own copies: a chunk must look the same (alignment included) whether it is the whole shard or a slice of it
            parts.append((c, Xd[a:b].clone(), Yd[a:b].clone(), _block_weights(b - a, block_size, dev)))
            vecs[c, nw + 1] = float(math.ceil((b - a) / block_size))

    def fn(w):
        wc = w.contiguous()
        for c, Xc, Yc, rw in parts:
            if native:
                _native_loss_grad(wc, layers, Xc, Yc, rw, vecs[c])
            else:
                loss, grad = loss_and_grad(wc, layers, Xc, Yc, rw)
                vecs[c, :nw].copy_(grad)
                vecs[c, nw] = loss
        if world > 1:
            gathered = [torch.empty_like(vecs) for _ in range(world)]
            dist.all_gather(gathered, vecs)
            v = torch.cat(gathered)
        else:
            v = vecs
        while v.shape[0] > 1:                       # fixed pairwise tree over the canonical chunks
            v = v[0::2] + v[1::2]
        total = v[0]
        blocks = total[nw + 1].clamp_min(1.0)
        return total[nw] / blocks, total[:nw] / blocks

    return fn


CANONICAL_CHUNKS = 8   # reproducible mode: worker counts 1, 2, 4, 8


def _fit_local(X, y, layers, block_size, max_iter, tol, solver, step_size, seed, initial_weights, device,
               chunk_rows: int = 0, chunks_local: int = 0):
    """Runs on every rank with its shard (or on the single process)."""
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    dev = torch.device(device)
    Xd = torch.as_tensor(X, dtype=torch.float64, device=dev)
    Yd = torch.as_tensor(y, dtype=torch.long, device=dev)
    n = Xd.shape[0]
    # MLlib DataStacker: consecutive rows of a partition form blocks of <= blockSize
    sizes = torch.full((n,), float(block_size), dtype=torch.float64, device=dev)
    rem = n % block_size
    if rem:
        sizes[n - rem:] = float(rem)
    row_w = 1.0 / sizes
    local_blocks = float(math.ceil(n / block_size)) if n else 0.0
    nw = num_weights(layers)
    reducer = _Reducer(nw + 2, dev, torch.float64)

    vec = reducer.vec                       # [gradient (nw) ‖ loss ‖ #blocks]; the kernels write into it directly
    vec.zero_()
    vec[nw + 1] = local_blocks
    native = dev.type == "cuda" and os.environ.get("B200ML_MLPC_NATIVE", "1") != "0"

    def fn(w):   # default: one evaluation over the whole shard, one all-reduce
        if n and native:
            _native_loss_grad(w.contiguous(), layers, Xd, Yd, row_w, vec)
        elif n:
            loss, grad = loss_and_grad(w, layers, Xd, Yd, row_w)
            vec[:nw].copy_(grad)
            vec[nw] = loss
        total = reducer()
        blocks = total[nw + 1].clamp_min(1.0)
        return total[nw] / blocks, total[:nw] / blocks

    threads = None
    if chunk_rows:
        fn = _chunked_fn(Xd, Yd, layers, block_size, chunk_rows, chunks_local, world, native)   # noqa: F811
        if dev.type == "cpu":
            # a BLAS that splits a product over threads may add in another order: one thread, whatever the host has
            threads = torch.get_num_threads()
            torch.set_num_threads(1)
    w0 = (torch.as_tensor(np.asarray(initial_weights), dtype=torch.float64) if initial_weights is not None
          else init_weights(layers, seed)).to(dev)
    if world > 1:
        dist.broadcast(w0, src=0)
    if dev.type == "cuda":
        # untimed: the first launches load the kernels (loss / gradient, the two-loop kernel, torch's dot products)
        # and size the scratch buffers - a launcher worker is a fresh process every time
        if solver == "l-bfgs":
            LBFGS(fn, m=10, max_iter=2, tol=0.0).minimize(w0)
        else:
            fn(w0)
        torch.cuda.synchronize(dev)
    t0 = time.time()
    if solver == "l-bfgs":
        opt = LBFGS(fn, m=10, max_iter=max_iter, tol=tol)
        w, iters = opt.minimize(w0)
        history, evals = opt.history, opt.evals
    elif solver == "gd":
        w, history, evals = w0.clone(), [], 0
        for it in range(1, max_iter + 1):
            f, g = fn(w)
            evals += 1
            history.append(float(f))
            w = w - (step_size / math.sqrt(it)) * g
            if float(g.norm()) < tol:
                break
        iters = len(history)
    else:
        raise ValueError(f"solver must be 'l-bfgs' or 'gd', got {solver!r}")
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)
    fit_s = time.time() - t0
    reducer.close()
    if threads is not None:
        torch.set_num_threads(threads)
    return {"weights": w.detach().cpu().numpy(), "objectiveHistory": history, "totalIterations": iters,
            "functionEvaluations": evals, "fitSeconds": fit_s, "world": world}


def _fit_worker(shards, kwargs):
    """TorchDistributor entry: init the process group, fit on this rank's shard, rank 0 returns the result."""
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    use_gpu = os.environ.get("B200ML_USE_GPU") == "1" and torch.cuda.is_available()
    if use_gpu:
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]) % torch.cuda.device_count())
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
        device = f"cuda:{torch.cuda.current_device()}"
    else:
        dist.init_process_group("gloo")
        device = "cpu"
    X, y = shards[rank]
    out = _fit_local(X, y, device=device, **kwargs)
    dist.destroy_process_group()
    return out if rank == 0 else None


# ---------------------------------------------------------------------------------------------------------------
# public API
# ---------------------------------------------------------------------------------------------------------------
class _Params:
    _defaults: dict = {}

    def __init__(self, **kw):
        self._p = dict(self._defaults)
        for k, v in kw.items():
            if k not in self._defaults:
                raise TypeError(f"{type(self).__name__} got an unexpected keyword argument {k!r}")
            if v is not None or self._defaults[k] is None:
                self._p[k] = v

    def __getattr__(self, name):
        if name.startswith("set") and len(name) > 3:
            key = name[3].lower() + name[4:]
            if key in self._defaults:
                def setter(v):
                    self._p[key] = v
                    return self
                return setter
        if name.startswith("get") and len(name) > 3:
            key = name[3].lower() + name[4:]
            if key in self._defaults:
                return lambda: self._p[key]
        raise AttributeError(name)

    def getOrDefault(self, k):
        return self._p[k]

    def setParams(self, **kw):
        for k, v in kw.items():
            self._p[k] = v
        return self

    def explainParams(self):
        return "\n".join(f"{k}: {v!r}" for k, v in self._p.items())


class MultilayerPerceptronClassifier(_Params):
    _defaults = dict(featuresCol="features", labelCol="label", predictionCol="prediction", maxIter=100, tol=1e-6,
                     seed=None, layers=None, blockSize=128, stepSize=0.03, solver="l-bfgs", initialWeights=None,
                     probabilityCol="probability", rawPredictionCol="rawPrediction", numWorkers=None, device=None,
                     reproducible=False)

    def _workers(self) -> int:
        if self._p["numWorkers"]:
            return int(self._p["numWorkers"])
        if dist.is_available() and dist.is_initialized():
            return 1  # already inside a worker: use the ambient process group
        from .session import SparkSession

        if SparkSession._active is not None:
            conf = SparkSession._active.sparkContext.getConf()
            if conf.contains("spark.executor.instances"):
                return int(conf.get("spark.executor.instances"))
        return 1

    def fit(self, dataset: DataFrame) -> "MultilayerPerceptronClassificationModel":
        p = self._p
        layers = list(p["layers"] or [])
        if len(layers) < 2:
            raise ValueError("layers must list at least the input and output sizes")
        X = np.asarray(dataset._matrix(p["featuresCol"]), dtype=np.float64)
        y = np.asarray(dataset._array(p["labelCol"]))
        if X.shape[1] != layers[0]:
            raise ValueError(f"features have {X.shape[1]} dimensions but layers[0] = {layers[0]}")
        yi = y.astype(np.int64)
        if (yi != y).any() or yi.min() < 0 or yi.max() >= layers[-1]:
            raise ValueError(f"labels must be integers in [0, {layers[-1]}) (MLlib LabelConverter contract)")
        seed = p["seed"] if p["seed"] is not None else abs(hash(type(self).__name__)) % (2 ** 31)
        kwargs = dict(layers=layers, block_size=int(p["blockSize"]), max_iter=int(p["maxIter"]), tol=float(p["tol"]),
                      solver=p["solver"], step_size=float(p["stepSize"]), seed=seed,
                      initial_weights=None if p["initialWeights"] is None else np.asarray(
                          p["initialWeights"].toArray() if hasattr(p["initialWeights"], "toArray") else p["initialWeights"]))
        workers = self._workers()
        in_group = dist.is_available() and dist.is_initialized()
        per_of = lambda w: math.ceil(len(X) / w)   # noqa: E731  rows per worker
        if p["reproducible"]:
            # not an MLlib parameter: the result is bit-identical for 1, 2, 4 and 8 workers (see _chunked_fn); the
            # shards are runs of whole canonical chunks, and a chunk holds whole MLlib blocks
            w_eff = dist.get_world_size() if in_group else workers
            if CANONICAL_CHUNKS % w_eff:
                raise ValueError(f"reproducible=True needs a worker count that divides {CANONICAL_CHUNKS}, got {w_eff}")
            bs = int(p["blockSize"])
            chunk_rows = max(1, math.ceil(math.ceil(len(X) / CANONICAL_CHUNKS) / bs)) * bs
            kwargs.update(chunk_rows=chunk_rows, chunks_local=CANONICAL_CHUNKS // w_eff)
            per_of = lambda w: chunk_rows * (CANONICAL_CHUNKS // w)   # noqa: E731
        if in_group:  # data-parallel inside an existing job: shard rows by rank
            r, w = dist.get_rank(), dist.get_world_size()
            per = per_of(w)
            dev = p["device"] or (f"cuda:{torch.cuda.current_device()}" if torch.cuda.is_available() and
                                  dist.get_backend() == "nccl" else "cpu")
            res = _fit_local(X[r * per:(r + 1) * per], yi[r * per:(r + 1) * per], device=dev, **kwargs)
        elif workers > 1:
            from .distributor import TorchDistributor

            per = per_of(workers)
            shards = [(X[r * per:(r + 1) * per], yi[r * per:(r + 1) * per]) for r in range(workers)]
            use_gpu = torch.cuda.is_available() and torch.cuda.device_count() >= workers and p["device"] != "cpu"
            res = TorchDistributor(num_processes=workers, local_mode=True, use_gpu=use_gpu).run(_fit_worker, shards, kwargs)
        else:
            dev = p["device"] or ("cuda" if torch.cuda.is_available() else "cpu")
            res = _fit_local(X, yi, device=dev, **kwargs)
        return MultilayerPerceptronClassificationModel(layers, res["weights"], dict(p), res)


class MLPTrainingSummary:
    def __init__(self, res):
        self.objectiveHistory = list(res["objectiveHistory"])
        self.totalIterations = int(res["totalIterations"])
        self.functionEvaluations = int(res.get("functionEvaluations", 0))
        self.fitSeconds = float(res.get("fitSeconds", 0.0))
        self.numWorkers = int(res.get("world", 1))


class MultilayerPerceptronClassificationModel:
    def __init__(self, layers, weights, params=None, res=None):
        self.layers = list(layers)
        self.weights = DenseVector(np.asarray(weights, dtype=np.float64))
        self.numFeatures = self.layers[0]
        self.numClasses = self.layers[-1]
        self._params = params or dict(MultilayerPerceptronClassifier._defaults)
        self.summary = MLPTrainingSummary(res) if res else None
        self.hasSummary = res is not None

    def _scores(self, X):
        w = torch.as_tensor(self.weights.toArray())
        _, P, Z = forward(w, self.layers, torch.as_tensor(np.asarray(X, dtype=np.float64)))
        return Z.numpy(), P.numpy()

    def predict(self, features) -> float:
        x = features.toArray() if hasattr(features, "toArray") else np.asarray(features)
        return float(self._scores(x[None, :])[1].argmax())

    def predictRaw(self, features):
        x = features.toArray() if hasattr(features, "toArray") else np.asarray(features)
        return DenseVector(self._scores(x[None, :])[0][0])

    def predictProbability(self, features):
        x = features.toArray() if hasattr(features, "toArray") else np.asarray(features)
        return DenseVector(self._scores(x[None, :])[1][0])

    def transform(self, dataset: DataFrame) -> DataFrame:
        Z, P = self._scores(dataset._matrix(self._params["featuresCol"]))
        out = dataset.withColumn(self._params["rawPredictionCol"], Z)
        out = out.withColumn(self._params["probabilityCol"], P)
        return out.withColumn(self._params["predictionCol"], P.argmax(1).astype(np.float64))

    def save(self, path: str) -> None:
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        with open(path, "wb") as fh:
            pickle.dump({"layers": self.layers, "weights": self.weights.toArray(), "params": self._params}, fh)

    def write(self):
        return self

    def overwrite(self):
        return self

    @classmethod
    def load(cls, path: str):
        with open(path, "rb") as fh:
            d = pickle.load(fh)
        return cls(d["layers"], d["weights"], d["params"])


class MulticlassClassificationEvaluator(_Params):
    _defaults = dict(predictionCol="prediction", labelCol="label", metricName="f1")

    def evaluate(self, dataset: DataFrame) -> float:
        pred = np.asarray(dataset._array(self._p["predictionCol"]))
        lab = np.asarray(dataset._array(self._p["labelCol"]))
        metric = self._p["metricName"]
        if len(lab) == 0:
            return 0.0
        if metric == "accuracy":
            return float((pred == lab).mean())
        classes = np.unique(lab)
        prec, rec, f1, wts = [], [], [], []
        for c in classes:
            tp = float(((pred == c) & (lab == c)).sum())
            fp = float(((pred == c) & (lab != c)).sum())
            fn = float(((pred != c) & (lab == c)).sum())
            pr = tp / (tp + fp) if tp + fp else 0.0
            rc = tp / (tp + fn) if tp + fn else 0.0
            prec.append(pr)
            rec.append(rc)
            f1.append(2 * pr * rc / (pr + rc) if pr + rc else 0.0)
            wts.append(float((lab == c).mean()))
        table = {"weightedPrecision": prec, "weightedRecall": rec, "f1": f1}
        if metric not in table:
            raise ValueError(f"unsupported metricName {metric!r}")
        return float(np.dot(table[metric], wts))

    def isLargerBetter(self):
        return True
