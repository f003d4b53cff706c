"""GEMM-shaped ops on the tcgen05 kernel (csrc/kernels/gemm_tcgen05.cu): raw ``gemm``, ``linear`` and the fused
multi-layer chain ``mlp_chain`` (Linear → act → dropout → Linear …) whose backward folds each activation's
derivative into the dgrad GEMM epilogue.

Reference behaviour: nn.Linear / nn.ReLU / nn.Sigmoid / nn.Dropout as used in transformer.py:71-72,104-117,177-179,272
and distributed_multilayer_perceptron.py:44-53.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native
from ..utils import workspace
from ..utils.rng import DeviceCounters

ACT_NONE, ACT_RELU, ACT_SIGMOID = 0, 1, 2
SPLITK_WORKSPACE_BYTES = 64 << 20
_ACT = {None: 0, "none": 0, "relu": 1, "sigmoid": 2, 0: 0, 1: 1, 2: 2}


def _ref_dropout_mask(shape, p, device):
    return (torch.rand(shape, device=device) >= p).float() / (1.0 - p)


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False, out: torch.Tensor | None = None,
         out_dtype=torch.bfloat16, bias: torch.Tensor | None = None, act=0, aux: torch.Tensor | None = None,
         aux_mode: int = 0, aux_scale: float = 1.0, alpha: float = 1.0, alpha_dev: torch.Tensor | None = None,
         accumulate: bool = False, dropout_p: float = 0.0, site: int = 0, tile_n: int = 0,
         force_simt: bool = False, n_logical: int | None = None, split_k: int = 0,
         epilogue_direct: bool = False, max_ctas: int = 0, row_stats: torch.Tensor | None = None) -> torch.Tensor:
    """C[M,N] = epilogue(alpha · A·Bᵀ).

    ``a`` is [M,K] (a_mn=False) or stored [K,M] (a_mn=True); ``b`` is [N,K] (b_mn=False) or stored [K,N]
    (b_mn=True).  Rows may be strided (stride(1) must be 1).  ``n_logical`` restricts N when ``b``/``out`` carry
    padding columns.  See include/b200ml.h for the epilogue order.
    """
    act = _ACT[act]
    assert a.dim() == 2 and b.dim() == 2 and a.stride(1) == 1 and b.stride(1) == 1
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    assert K == Kb, f"reduction mismatch {K} vs {Kb}"
    if n_logical is not None:
        N = n_logical
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    assert out.stride(1) == 1 and out.shape[0] >= M and out.shape[1] >= N

    if not _native.use_native(a):
        # PyTorch restatement (CPU): same math in fp32
        A = (a.t() if a_mn else a).float()
        B = (b.t() if b_mn else b).float()[:N]
        v = alpha * (A @ B.t())
        if alpha_dev is not None:
            v = v * alpha_dev.float()
        if bias is not None:
            v = v + bias.float()[:N]
        if act == ACT_RELU:
            v = torch.relu(v)
        elif act == ACT_SIGMOID:
            v = torch.sigmoid(v)
        if dropout_p > 0:
            v = v * _ref_dropout_mask(v.shape, dropout_p, v.device)
        if aux is not None and aux_mode:
            ax = aux[:M, :N].float()
            if aux_mode == 1:
                v = torch.where(ax > 0, v * aux_scale, torch.zeros_like(v))
            elif aux_mode == 2:
                v = v + ax
            else:
                v = v * ax * (1 - ax)
        tgt = out[:M, :N]
        if accumulate:
            tgt.add_(v.to(out.dtype))
        else:
            tgt.copy_(v.to(out.dtype))
        return out

    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16, "native gemm takes bf16 operands"
    assert out.dtype in (torch.bfloat16, torch.float32)
    lib = _native.load()
    g = _native.GemmArgs()
    g.M, g.N, g.K = M, N, K
    g.A, g.lda, g.a_mn_major = a.data_ptr(), a.stride(0), int(a_mn)
    g.B, g.ldb, g.b_mn_major = b.data_ptr(), b.stride(0), int(b_mn)
    g.C, g.ldc, g.c_fp32 = out.data_ptr(), out.stride(0), int(out.dtype == torch.float32)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.is_contiguous()
        g.bias = bias.data_ptr()
    g.act = act
    if aux is not None and aux_mode:
        assert aux.dtype == torch.bfloat16 and aux.stride(1) == 1
        g.aux, g.ldaux, g.aux_mode, g.aux_scale = aux.data_ptr(), aux.stride(0), aux_mode, aux_scale
    g.alpha = alpha
    if alpha_dev is not None:
        assert alpha_dev.dtype == torch.float32
        g.alpha_dev = alpha_dev.data_ptr()
    g.accumulate = int(accumulate)
    if dropout_p > 0:
        g.dropout_p = dropout_p
        g.rng_state = DeviceCounters.get(a.device).rng.data_ptr()
        g.rng_site = site
    g.tile_n = tile_n
    g.force_simt = int(force_simt)
    g.split_k = split_k
    g.epilogue_direct = int(epilogue_direct)
    g.max_ctas = int(max_ctas)
    if row_stats is not None:   # float32 [M, P, 2]: per-row (max, sum-exp) of every half tile (include/b200ml.h)
        assert row_stats.dtype == torch.float32 and row_stats.is_contiguous() and row_stats.shape[0] >= M
        g.row_stats, g.row_stats_ld = row_stats.data_ptr(), row_stats.shape[1]
    if split_k != 1 and bias is None and not act and not (aux is not None and aux_mode) and dropout_p <= 0:
        ws = workspace.get(f"gemm_splitk@{workspace.lane()}", SPLITK_WORKSPACE_BYTES, a.device)  # one per stream lane
        g.workspace, g.workspace_bytes = ws.data_ptr(), ws.numel()
    _native.check(lib.b200ml_gemm_bf16(C.byref(g), _native.stream_ptr()), "gemm_bf16")
    return out


def colsum(x: torch.Tensor, out: torch.Tensor | None = None, accumulate: bool = False) -> torch.Tensor:
    """Column sums of a bf16 [rows, cols] matrix in fp32 (bias gradients); deterministic two-stage reduction."""
    rows, cols = x.shape
    if out is None:
        out = torch.empty(cols, dtype=torch.float32, device=x.device)
    if not _native.use_native(x):
        s = x.float().sum(0)
        out.add_(s) if accumulate else out.copy_(s)
        return out
    lib = _native.load()
    nbytes = lib.b200ml_colsum_scratch_bytes(rows, cols)
    ws = workspace.get(f"colsum@{workspace.lane()}", nbytes, x.device)  # one per stream lane
    _native.check(lib.b200ml_colsum_bf16(x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), int(accumulate),
                                         ws.data_ptr(), ws.numel(), _native.stream_ptr()), "colsum")
    return out


# ----------------------------------------------------------------------------------------------------------------
# parameter plumbing shared by every op with weights
# ----------------------------------------------------------------------------------------------------------------
def working_copy(param: torch.Tensor) -> torch.Tensor:
    """bf16 working copy of an fp32 master parameter.

    The DP engine / fused optimizer refreshes it in the same kernel that updates the master weights
    (``param._b200_w16`` is then a view into the flat bf16 buffer).  Without an engine the copy is re-cast when the
    parameter's version counter moves.
    """
    if param.dtype == torch.bfloat16 or not param.is_cuda:
        return param.detach()  # CPU restatement computes in the parameter's own precision
    w16 = getattr(param, "_b200_w16", None)
    if w16 is not None and getattr(param, "_b200_w16_managed", False):
        return w16
    ver = param._version
    if w16 is None or getattr(param, "_b200_w16_ver", None) != ver or w16.device != param.device:
        w16 = param.detach().to(torch.bfloat16)
        param._b200_w16 = w16
        param._b200_w16_ver = ver
    return w16


class GradSideStream:
    """Parameter-gradient kernels (wgrad GEMMs, bias column sums) are leaves of the backward graph: nothing else in
    backward consumes them.  The engine lets them run on side streams so they overlap the dgrad / attention /
    LayerNorm chain (small grids leave SMs idle); inside a captured CUDA graph the streams become parallel branches.
    Several streams are used round-robin (``B200ML_WGRAD_STREAMS``, default 3): a single one serialises the step's
    eleven small weight-gradient chains (GEMM + column sum + partial reduce, ~40 us each, a handful of CTAs each) and
    the measured timeline showed that chain finishing ~150 us after the rest of backward.  Every stream owns its
    scratch lane.  Tensors the side kernels read are kept alive here until the engine has joined the streams."""

    def __init__(self, device, n_streams: int | None = None):
        import os

        n = n_streams if n_streams is not None else int(os.environ.get("B200ML_WGRAD_STREAMS", "3"))
        self.streams = [torch.cuda.Stream(device=device) for _ in range(max(n, 1))]
        for k, s in enumerate(self.streams):
            workspace.register_stream(s, 10 + k)
        self.stream = self.streams[0]
        self.last = self.streams[0]   # the stream the most recent run() used (the engine records its event there)
        self._next = 0
        self._touched: set = set()
        self.keep: list = []
        self.active = False
        self.used = False  # has work been queued since the last join (i.e. are the streams part of the current step/capture)

    def run(self, fn, *args) -> None:
        cur = torch.cuda.current_stream()
        k = self._next
        self._next = (k + 1) % len(self.streams)
        st = self.streams[k]
        st.wait_stream(cur)
        with torch.cuda.stream(st):
            fn(*args)
        self.last = st
        self._touched.add(k)
        self.used = True
        self.keep.append(fn)  # the closure owns references to the gradient / activation tensors it reads

    def join(self) -> None:
        if self.used:
            cur = torch.cuda.current_stream()
            for k in sorted(self._touched):
                cur.wait_stream(self.streams[k])
        self.used = False
        self._touched.clear()
        self._next = 0
        self.keep.clear()


def deliver_grad(param: torch.Tensor, compute, shape) -> torch.Tensor | None:
    """Produce a parameter gradient.  With a DP engine the gradient is written straight into the parameter's slot of
    the flat bucket (``param._b200_grad``) and the engine is told the slot is ready — autograd sees ``None``.
    Otherwise a fresh fp32 tensor is returned to autograd."""
    slot = getattr(param, "_b200_grad", None)
    if slot is not None and slot.device == param.device:
        acc = bool(getattr(param, "_b200_grad_accumulate", False))
        side = getattr(param, "_b200_side", None)
        if side is not None and side.active and slot.is_cuda:
            side.run(compute, slot.view(shape), acc)
        else:
            compute(slot.view(shape), acc)
        ready = getattr(param, "_b200_ready", None)
        if ready is not None:
            ready()
        return None
    out = torch.empty(shape, dtype=torch.float32, device=param.device)
    compute(out, False)
    return out.to(param.dtype) if param.dtype != torch.float32 else out


# ----------------------------------------------------------------------------------------------------------------
# fused Linear chain
# ----------------------------------------------------------------------------------------------------------------
class _MLPChain(torch.autograd.Function):
    """y = L_n(... act_1(L_1(x)))   with per-layer activation ∈ {none, relu, sigmoid} and optional dropout after it.

    Saves only the layer inputs (bf16).  Backward per layer: dW = dZᵀ·X, db = colsum(dZ), dX = dZ·W with the previous
    layer's activation derivative applied in the dgrad epilogue (aux_mode 1 / 3), so no elementwise pass exists.

    ``fork``: the op also returns its input as a second output, to be used as the residual branch of the LayerNorm
    that closes the sub-layer (y = LN(chain(x) + x)).  Autograd then hands backward the residual branch's gradient
    separately instead of summing it with an elementwise kernel, and the first layer's dgrad GEMM adds it in its
    epilogue (aux_mode 2): dX = dZ·W + d_skip.  ``skip_last_bias``: the last layer's bias gradient is produced elsewhere
    (by the LayerNorm backward that consumes this op's output, ops/norm.py)."""

    @staticmethod
    def forward(ctx, x, n_layers, acts, drops, sites, fork, skip_last_bias, *params):
        weights = params[:n_layers]
        biases = params[n_layers:]
        lead = x.shape[:-1]
        h = x.reshape(-1, x.shape[-1])
        native = _native.use_native(x)
        if native and h.dtype != torch.bfloat16:
            h = h.to(torch.bfloat16)
        if not h.is_contiguous():
            h = h.contiguous()
        inputs = []
        for i in range(n_layers):
            inputs.append(h)
            w16 = working_copy(weights[i])
            b = biases[i]
            h = gemm(h, w16, bias=None if b is None else b.detach(), act=acts[i], dropout_p=drops[i], site=sites[i],
                     out_dtype=torch.bfloat16 if native else h.dtype)
        ctx.n_layers, ctx.acts, ctx.drops = n_layers, acts, drops
        ctx.fork, ctx.skip_last_bias = fork, skip_last_bias
        ctx.params = params
        ctx.x_dtype = x.dtype
        ctx.save_for_backward(*inputs, h)
        out = h.reshape(*lead, h.shape[-1])
        return (out, x) if fork else out

    @staticmethod
    def backward(ctx, dy, dskip=None):
        n = ctx.n_layers
        weights, biases = ctx.params[:n], ctx.params[n:]
        saved = ctx.saved_tensors
        inputs, y = saved[:n], saved[n]
        if dy is None:   # only the skip output was used
            return (dskip, None, None, None, None, None, None, *([None] * (2 * n)))
        dz = dy.reshape(-1, dy.shape[-1])
        native = _native.use_native(dz)
        cdt = torch.bfloat16 if native else y.dtype
        if dz.dtype != cdt:
            dz = dz.to(cdt)
        if not dz.is_contiguous():
            dz = dz.contiguous()
        # last layer's own activation (rare: chain ending in an activation)
        if ctx.acts[n - 1]:
            yf = y.float()
            if ctx.acts[n - 1] == ACT_RELU:
                dz = (dz.float() * (yf > 0) / (1.0 - ctx.drops[n - 1])).to(cdt)
            else:
                assert ctx.drops[n - 1] == 0
                dz = (dz.float() * yf * (1 - yf)).to(cdt)
        gw = [None] * n
        gb = [None] * n
        dx = None
        for i in range(n - 1, -1, -1):
            xin = inputs[i]
            w = weights[i]
            N, K = w.shape
            # dgrad FIRST: delivering a parameter gradient may complete a bucket, and the engine then updates that
            # bucket's weights on its side stream — every read of this layer's weights must already be enqueued.
            dz_this = dz
            need_dx = i > 0 or ctx.needs_input_grad[0]
            if need_dx:
                w16 = working_copy(w)
                if i > 0 and ctx.acts[i - 1] == ACT_RELU:
                    dz = gemm(dz, w16, b_mn=True, aux=xin, aux_mode=1, aux_scale=1.0 / (1.0 - ctx.drops[i - 1]),
                              out_dtype=cdt)
                elif i > 0 and ctx.acts[i - 1] == ACT_SIGMOID:
                    assert ctx.drops[i - 1] == 0, "sigmoid followed by dropout is not fused"
                    dz = gemm(dz, w16, b_mn=True, aux=xin, aux_mode=3, out_dtype=cdt)
                elif i == 0 and dskip is not None:
                    sk = dskip.reshape(-1, K)
                    if sk.dtype != cdt:
                        sk = sk.to(cdt)
                    if not sk.is_contiguous():
                        sk = sk.contiguous()
                    dz = gemm(dz, w16, b_mn=True, aux=sk, aux_mode=2, out_dtype=cdt)   # dX = dZ·W + d_skip
                else:
                    dz = gemm(dz, w16, b_mn=True, out_dtype=cdt)
                if i == 0:
                    dx = dz
            if w.requires_grad:
                gw[i] = deliver_grad(w, lambda out, acc, g=dz_this, xin=xin: gemm(g, xin, a_mn=True, b_mn=True, out=out,
                                                                                 accumulate=acc), (N, K))
            b = biases[i]
            if b is not None and b.requires_grad and not (ctx.skip_last_bias and i == n - 1):
                gb[i] = deliver_grad(b, lambda out, acc, g=dz_this: colsum(g, out=out, accumulate=acc), (N,))
        if dx is not None:
            dx = dx.reshape(*dy.shape[:-1], dx.shape[-1]).to(ctx.x_dtype)
        elif dskip is not None and ctx.needs_input_grad[0]:
            dx = dskip
        return (dx, None, None, None, None, None, None, *gw, *gb)


def mlp_chain(x, weights, biases, acts, drops=None, sites=None, training: bool = True, fork: bool = False,
              skip_last_bias_grad: bool = False):
    """Fused chain of Linear layers; ``acts[i]`` follows layer i ('relu' | 'sigmoid' | None), ``drops[i]`` follows it.
    ``fork=True`` returns ``(y, x_skip)``; see _MLPChain."""
    n = len(weights)
    acts = tuple(_ACT[a] for a in acts)
    drops = tuple((d if training else 0.0) for d in (drops or [0.0] * n))
    sites = tuple(sites or [0] * n)
    return _MLPChain.apply(x, n, acts, drops, sites, bool(fork), bool(skip_last_bias_grad), *weights, *biases)


def linear(x, weight, bias=None, act=None, dropout_p: float = 0.0, site: int = 0, training: bool = True,
           fork: bool = False, skip_bias_grad: bool = False):
    """y = dropout(act(x·Wᵀ + b)) — nn.Linear with its activation/dropout fused into the GEMM epilogue."""
    return mlp_chain(x, [weight], [bias], [act], [dropout_p], [site], training, fork, skip_bias_grad)
