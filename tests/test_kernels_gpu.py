"""LayerNorm / embedding / cross-entropy / optimizer kernels vs PyTorch fp32 restatements."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("rows,d", [(64, 512), (6400, 512), (37, 64), (100, 1024)])
@pytest.mark.parametrize("with_res", [True, False])
def test_layernorm_fwd_bwd(rows, d, with_res):
    from b200ml.ops.norm import layernorm_residual
    torch.manual_seed(0)
    a = torch.randn(rows, d, device="cuda").to(torch.bfloat16).requires_grad_()
    x = torch.randn(rows, d, device="cuda").to(torch.bfloat16).requires_grad_() if with_res else None
    gamma = (1 + 0.1 * torch.randn(d, device="cuda")).requires_grad_()
    beta = (0.1 * torch.randn(d, device="cuda")).requires_grad_()
    y = layernorm_residual(a, x, gamma, beta)
    w = torch.randn(rows, d, device="cuda")
    (y.float() * w).sum().backward()

    ar = a.detach().float().cpu().requires_grad_()
    xr = x.detach().float().cpu().requires_grad_() if with_res else None
    gr, br = gamma.detach().cpu().requires_grad_(), beta.detach().cpu().requires_grad_()
    z = ar + xr if with_res else ar
    mean = z.mean(-1, keepdim=True)
    var = ((z - mean) ** 2).mean(-1, keepdim=True)
    yr = gr * ((z - mean) / (var + 1e-5).sqrt()) + br
    (yr * w.cpu()).sum().backward()
    assert _rel(y, yr) < 0.01
    assert _rel(a.grad, ar.grad) < 0.02
    if with_res:
        assert _rel(x.grad, xr.grad) < 0.02
    assert _rel(gamma.grad, gr.grad) < 0.02
    assert _rel(beta.grad, br.grad) < 0.02


def test_layernorm_dropout_consistent():
    """Backward regenerates the forward's dropout mask: da is exactly zero where the forward dropped the element and
    equals dz / (1 - p) where it kept it."""
    from b200ml.ops.norm import layernorm_residual
    from b200ml.utils.rng import DeviceCounters
    DeviceCounters.get("cuda").seed(5)
    rows, d, p = 256, 512, 0.5
    # a = 1 everywhere and no residual: the pre-norm value is 0 (dropped) or 1/(1-p) (kept), which exposes the mask
    a = torch.ones(rows, d, device="cuda", dtype=torch.bfloat16).requires_grad_()
    g = (torch.rand(d, device="cuda") + 0.5).requires_grad_()
    b = torch.zeros(d, device="cuda", requires_grad=True)
    y = layernorm_residual(a, None, g, b, dropout_p=p, site=3)
    kept_fwd = y.detach().float() > 0        # kept elements sit above the row mean, dropped ones below (gamma > 0)
    frac = 1.0 - kept_fwd.float().mean().item()
    assert 0.45 < frac < 0.55
    w = torch.randn(rows, d, device="cuda")
    (y.float() * w).sum().backward()
    assert torch.equal(a.grad != 0, kept_fwd) or ((a.grad != 0) ^ kept_fwd).float().mean().item() < 1e-3
    # with a residual input the same dz reaches both inputs: da == dx / (1 - p) where kept, 0 where dropped
    a2 = torch.ones(rows, d, device="cuda", dtype=torch.bfloat16).requires_grad_()
    x2 = torch.zeros(rows, d, device="cuda", dtype=torch.bfloat16).requires_grad_()
    y2 = layernorm_residual(a2, x2, g, b, dropout_p=p, site=3)
    assert torch.equal(y2, y)                # same (seed, step, site): same draw
    (y2.float() * w).sum().backward()
    da, dx = a2.grad.float(), x2.grad.float()
    assert torch.equal(da[~kept_fwd], torch.zeros_like(da[~kept_fwd]))
    want = dx[kept_fwd] / (1 - p)
    assert (da[kept_fwd] - want).abs().max().item() <= 1e-2 * want.abs().max().item() + 1e-6


def test_dropout_masks_survive_counter_bumps_between_forward_and_backward():
    """The reference's LSTM loop calls zero_grad() between forward and backward (distributed_lstm.py:192-194).  The
    engine's begin_step() must not move the dropout draw: gradients are the same wherever zero_grad() sits."""
    from b200ml.models import LSTM
    from b200ml.ops.loss import cross_entropy
    from b200ml.parallel.engine import DataParallelEngine
    from b200ml.utils.rng import manual_seed
    tok = torch.randint(1, 400, (8, 24), device="cuda")
    lab = torch.randint(0, 4, (8,), device="cuda")
    z = torch.zeros(2, 8, 32, device="cuda")
    grads = []
    import itertools
    from b200ml.utils import rng as rng_mod
    for order in ("before", "between"):
        manual_seed(3, "cuda")
        rng_mod._site_counter = itertools.count(5000)   # both models get the same dropout site ids
        model = LSTM(400, 32, 32, 4).cuda().train()
        eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)
        for _ in range(2):                   # second step: fresh masks, still consistent
            if order == "before":
                eng.zero_grad()
            pred, _, _ = model(tok, z, z)
            loss = cross_entropy(pred[:, -1, :], lab)
            if order == "between":
                eng.zero_grad()
            loss.backward()
            torch.cuda.synchronize()
            grads.append(eng.flat.flat_g.clone())
            eng.step()
        eng.close()
    for a, b in ((grads[0], grads[2]), (grads[1], grads[3])):
        assert torch.equal(a, b)             # every kernel of the step is deterministic
    assert not torch.allclose(grads[0], grads[1])


def test_flat_optimizer_advances_dropout_step():
    """Two consecutive FlatOptimizer steps must draw different dropout masks (pytorch_lstm.py trains without an engine)."""
    from b200ml.ops.embedding import embedding
    from b200ml.ops.optim import FlatOptimizer
    from b200ml.utils.rng import manual_seed
    manual_seed(9, "cuda")
    w = torch.nn.Parameter(torch.ones(64, 128, device="cuda"))
    opt = FlatOptimizer([w], kind="sgd", lr=0.0)
    tok = torch.arange(64, device="cuda")[None]
    masks = []
    for _ in range(2):
        opt.zero_grad()
        out = embedding(tok, w, dropout_p=0.5, site=77, out_dtype=torch.float32)
        out.sum().backward()
        masks.append(out.detach() != 0)
        # backward saw the forward's draw: the gradient rows are non-zero exactly where the output was kept
        assert torch.equal(w._b200_grad.view(64, 128) != 0, masks[-1][0])
        opt.step()
    assert not torch.equal(masks[0], masks[1])


def test_embedding_pe_matches_reference_formula():
    from b200ml.ops.embedding import embedding, positional_encoding_table
    torch.manual_seed(0)
    V, d, B, S = 1000, 512, 4, 200
    w = torch.randn(V, d, device="cuda").requires_grad_()
    tok = torch.randint(0, V, (B, S), device="cuda")
    pe = positional_encoding_table(S, d, "cuda")
    pe_cpu = positional_encoding_table(S, d, "cpu")
    assert (pe.cpu() - pe_cpu).abs().max().item() < 2e-3
    out = embedding(tok, w, pe=pe)
    ref = w.detach().to(torch.bfloat16).float().cpu()[tok.cpu()] + pe_cpu
    assert _rel(out, ref) < 0.01
    gout = torch.randn(B, S, d, device="cuda")
    (out.float() * gout).sum().backward()
    gref = torch.zeros(V, d).index_add_(0, tok.cpu().reshape(-1), gout.cpu().to(torch.bfloat16).float().reshape(-1, d))
    assert _rel(w.grad, gref) < 0.01


def test_embedding_padding_idx():
    from b200ml.ops.embedding import embedding
    V, d = 50, 32
    w = torch.randn(V, d, device="cuda").requires_grad_()
    tok = torch.tensor([[0, 3, 0, 7]], device="cuda")
    out = embedding(tok, w, padding_idx=0, out_dtype=torch.float32)
    out.sum().backward()
    assert w.grad[0].abs().sum().item() == 0
    assert w.grad[3].sum().item() == pytest.approx(d)


@pytest.mark.parametrize("n_tokens,V,d", [(6400, 19214, 512), (4128, 95815, 32), (37, 11, 64), (16384, 300, 64),
                                          (20000, 300, 64)])
def test_embedding_backward_is_deterministic_and_exact(n_tokens, V, d):
    """Sort + segment reduction: every table row written once, fp32 sums in position order -> identical bits on every
    run, equal to the fp32 index_add_ oracle up to summation order; untouched rows and the padding row are exactly 0
    even when the gradient buffer held garbage.  20000 tokens exceed the single-CTA sort and take the atomicAdd path
    (same values up to summation order, not bit-reproducible)."""
    from b200ml.ops.embedding import embedding
    torch.manual_seed(1)
    tok = torch.randint(0, V, (1, n_tokens), device="cuda")
    tok[0, ::7] = 3                                      # a heavily repeated token
    tok[0, 1::11] = 0                                    # padding
    gout = torch.randn(1, n_tokens, d, device="cuda")
    grads = []
    for _ in range(3):
        w = torch.randn(V, d, device="cuda").requires_grad_()
        out = embedding(tok, w, padding_idx=0, out_dtype=torch.float32)
        (out * gout).sum().backward()
        grads.append(w.grad.clone())
    if n_tokens <= 16384:
        assert torch.equal(grads[0], grads[1]) and torch.equal(grads[0], grads[2])
    g = gout.reshape(-1, d).clone()
    g[tok.reshape(-1) == 0] = 0
    ref = torch.zeros(V, d, device="cuda", dtype=torch.float64).index_add_(0, tok.reshape(-1), g.double())
    assert (grads[0].double() - ref).abs().max().item() <= 1e-5 * ref.abs().max().item()
    assert grads[0][0].abs().max().item() == 0
    untouched = torch.ones(V, dtype=torch.bool, device="cuda")
    untouched[tok.reshape(-1)] = False
    assert grads[0][untouched].abs().max().item() == 0 if untouched.any() else True
    # straight into a dirty flat gradient slot: still overwritten, never accumulated
    from b200ml.ops.optim import FlatOptimizer
    w = torch.nn.Parameter(torch.randn(V, d, device="cuda"))
    opt = FlatOptimizer([w], kind="sgd", lr=0.0)
    w._b200_grad.fill_(123.0)
    out = embedding(tok, w, padding_idx=0, out_dtype=torch.float32)
    (out * gout).sum().backward()
    if n_tokens <= 16384:
        assert torch.equal(w._b200_grad.view(V, d), grads[0])
    else:
        assert (w._b200_grad.view(V, d) - grads[0]).abs().max().item() <= 1e-5 * grads[0].abs().max().item()


@pytest.mark.parametrize("rows,V,dtype", [(32, 10, torch.float32), (30, 3, torch.float32), (640, 19214, torch.bfloat16)])
def test_cross_entropy(rows, V, dtype):
    from b200ml.ops.loss import cross_entropy
    torch.manual_seed(0)
    z = (torch.randn(rows, V, device="cuda") * 2).to(dtype).requires_grad_()
    t = torch.randint(0, V, (rows,), device="cuda")
    t[::5] = 0
    loss = cross_entropy(z, t, ignore_index=0)
    loss.backward()
    zr = z.detach().float().cpu().requires_grad_()
    lr = torch.nn.functional.cross_entropy(zr, t.cpu(), ignore_index=0, reduction="none")
    valid = t.cpu() != 0
    lref = lr[valid].sum() / valid.sum()
    lref.backward()
    assert abs(loss.item() - lref.item()) < 2e-3 * max(1.0, abs(lref.item()))
    assert _rel(z.grad, zr.grad) < (0.02 if dtype == torch.bfloat16 else 1e-4)


@pytest.mark.parametrize("rows,V", [(512, 19214), (77, 1000), (300, 20480), (64, 20488)])
def test_linear_cross_entropy_head(rows, V):
    """V <= 20480 takes the persistent row-cached loss kernel with the fused bias-gradient column sums (several rows
    per CTA when rows > SMs, ragged last vector when V % 8 != 0); V = 20488 falls back to the generic kernel."""
    from b200ml.ops.loss import linear_cross_entropy
    torch.manual_seed(0)
    d = 512
    h = (torch.randn(rows, d, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_()
    w = (torch.randn(V, d, device="cuda") * 0.05).requires_grad_()
    b = (torch.randn(V, device="cuda") * 0.05).requires_grad_()
    t = torch.randint(0, V, (rows,), device="cuda")
    t[::3] = 0
    loss = linear_cross_entropy(h, w, b, t, ignore_index=0)
    (loss * 2.0).backward()
    hr = h.detach().float().cpu().requires_grad_()
    wr = w.detach().to(torch.bfloat16).float().cpu().requires_grad_()
    br = b.detach().cpu().requires_grad_()
    lr = torch.nn.functional.cross_entropy(hr @ wr.t() + br, t.cpu(), ignore_index=0, reduction="none")
    valid = t.cpu() != 0
    lref = lr[valid].sum() / valid.sum()
    (lref * 2.0).backward()
    assert abs(loss.item() - lref.item()) < 5e-3 * abs(lref.item())
    assert _rel(h.grad, hr.grad) < 0.03
    assert _rel(w.grad, wr.grad) < 0.03
    assert _rel(b.grad, br.grad) < 0.03


def test_head_statistics_from_gemm_epilogue_match_row_kernel(monkeypatch):
    """The head GEMM's epilogue emits per-row (max, sum-exp) pairs and the loss kernel makes one pass; the result must
    be the one of the self-contained three-pass kernel (same bf16 logits, same statistics up to fp32 rounding), also
    with a ragged last tile (V % 256 != 0, rows % 256 != 0) and ignored rows."""
    from b200ml.ops import loss as L
    torch.manual_seed(1)
    rows, V, d = 700, 19214, 512
    h0 = (torch.randn(rows, d, device="cuda") * 0.5).to(torch.bfloat16)
    w0 = torch.randn(V, d, device="cuda") * 0.05
    b0 = torch.randn(V, device="cuda") * 0.05
    t = torch.randint(0, V, (rows,), device="cuda")
    t[::4] = 0
    res = []
    for stats in (True, False):
        monkeypatch.setattr(L, "HEAD_STATS", stats)
        h, w, b = h0.clone().requires_grad_(), w0.clone().requires_grad_(), b0.clone().requires_grad_()
        loss = L.linear_cross_entropy(h, w, b, t, ignore_index=0)
        loss.backward()
        res.append((loss.detach(), h.grad, w.grad, b.grad))
    assert abs(res[0][0].item() - res[1][0].item()) < 1e-5 * abs(res[1][0].item())
    for a, c in zip(res[0][1:], res[1][1:]):
        assert _rel(a, c) < 2e-3


@pytest.mark.parametrize("kind", ["sgd", "adam"])
def test_flat_optimizer_matches_torch(kind):
    from b200ml.ops.optim import FlatOptimizer
    torch.manual_seed(0)
    ps = [torch.randn(257, 33, device="cuda").requires_grad_(), torch.randn(1000, device="cuda").requires_grad_()]
    rs = [p.detach().clone().requires_grad_() for p in ps]
    opt = FlatOptimizer(ps, kind=kind, lr=0.05)
    ref = (torch.optim.SGD if kind == "sgd" else torch.optim.Adam)(rs, lr=0.05)
    for step in range(3):
        for p, r in zip(ps, rs):
            g = torch.randn_like(p)
            p.grad = g.clone()
            r.grad = g.clone()
        opt.step()
        ref.step()
    for p, r in zip(ps, rs):
        assert (p - r).abs().max().item() < 1e-5
        assert (p._b200_w16.float() - p.detach()).abs().max().item() < 0.02 * p.abs().max().item()


@pytest.mark.parametrize("n,k", [(64, 0), (64, 3), (5000, 10), (535818, 10)])
def test_lbfgs_direction_kernel_matches_two_loop(n, k):
    """K20: cooperative two-loop kernel vs the plain torch recursion (float64), and run-to-run bit reproducibility."""
    from b200ml.spark.ml import LBFGS, _native_direction
    torch.manual_seed(0)
    g = torch.randn(n, dtype=torch.float64, device="cuda")
    S = [torch.randn(n, dtype=torch.float64, device="cuda") * 0.1 for _ in range(k)]
    Y = [s * (1.0 + 0.5 * torch.rand(n, dtype=torch.float64, device="cuda")) for s in S]   # <s, y> > 0
    d = _native_direction(g, S, Y)
    assert d is not None
    opt = LBFGS(lambda w: None)
    ref = LBFGS.two_loop(opt, g.cpu(), [s.cpu() for s in S], [y.cpu() for y in Y])   # CPU tensors -> torch path
    assert ((d.cpu() - ref).norm() / ref.norm()).item() < 1e-11
    assert torch.equal(d, _native_direction(g, S, Y))


@pytest.mark.parametrize("layers,rows,block", [([4, 5, 4, 3], 150, 128), ([784, 64, 10], 333, 128), ([20, 130, 70, 65, 7], 1000, 30)])
def test_mlpc_native_loss_grad_matches_float64_restatement(layers, rows, block):
    """K19: the float64 GEMM / softmax kernels working on MLlib's flat weight layout against the torch float64
    restatement of the same math (spark/ml.py loss_and_grad), including the per-block row weights."""
    from b200ml.spark import ml
    torch.manual_seed(0)
    dev = torch.device("cuda")
    X = torch.randn(rows, layers[0], dtype=torch.float64, device=dev)
    Y = torch.randint(0, layers[-1], (rows,), device=dev)
    sizes = torch.full((rows,), float(block), dtype=torch.float64, device=dev)
    if rows % block:
        sizes[rows - rows % block:] = float(rows % block)
    row_w = 1.0 / sizes
    w = ml.init_weights(layers, 7).to(dev)
    nw = ml.num_weights(layers)
    loss_ref, grad_ref = ml.loss_and_grad(w, layers, X, Y, row_w)
    outs = []
    for _ in range(2):
        vec = torch.full((nw + 2,), 123.0, dtype=torch.float64, device=dev)
        ml._native_loss_grad(w, layers, X, Y, row_w, vec)
        outs.append(vec.clone())
    assert torch.equal(outs[0], outs[1])                    # deterministic
    vec = outs[0]
    assert vec[nw + 1].item() == 123.0                      # the block count slot is the caller's
    assert abs(vec[nw].item() - loss_ref.item()) <= 1e-12 * max(1.0, abs(loss_ref.item()))
    assert (vec[:nw] - grad_ref).abs().max().item() <= 1e-12 * max(1.0, grad_ref.abs().max().item())


def test_mlpc_fit_on_gpu_uses_native_kernels_and_matches_cpu(monkeypatch):
    from b200ml import _native
    from b200ml.spark import ml
    rng = torch.Generator().manual_seed(3)
    X = torch.randn(300, 4, generator=rng, dtype=torch.float64).numpy()
    y = (torch.randn(300, generator=rng) > 0).long().numpy() + (X[:, 0] > 0.5)
    kw = dict(layers=[4, 5, 4, 3], block_size=128, max_iter=15, tol=1e-9, solver="l-bfgs", step_size=0.03, seed=11,
              initial_weights=None)
    n0 = _native.launches()
    g = ml._fit_local(X, y, device="cuda", **kw)
    assert _native.launches() - n0 >= 15                     # every evaluation goes through this library's kernels
    c = ml._fit_local(X, y, device="cpu", **kw)
    assert abs(g["objectiveHistory"][-1] - c["objectiveHistory"][-1]) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("rows,C,dtype", [(32, 10, torch.float32), (257, 4, torch.float32), (64, 19214, torch.bfloat16),
                                          (5, 3, torch.bfloat16)])
def test_argmax_accuracy_kernel_matches_torch(rows, C, dtype):
    """K14 (distributed_cnn.py:144-147,211-212): predicted class = arg-max of the logits (softmax is monotone), the
    device counter accumulates the number of correct rows over calls; first maximum wins on ties, like torch.argmax."""
    from b200ml.ops.metrics import argmax_accuracy
    torch.manual_seed(rows * 31 + C)
    z = torch.randn(rows, C, device="cuda").to(dtype)
    z[0, :] = 1.0                                      # a full tie: class 0
    if C > 2:
        z[1, 1] = z[1, 2] = 50.0                       # a two-way tie: the lower index
    y = torch.randint(0, C, (rows,), device="cuda")
    pred, correct = argmax_accuracy(z, y)
    ref = z.float().argmax(1)
    assert torch.equal(pred, ref)
    assert int(correct.item()) == int((ref == y).sum())
    pred2, correct = argmax_accuracy(z, y, correct)    # accumulates
    assert int(correct.item()) == 2 * int((ref == y).sum())
    pred3, c3 = argmax_accuracy(z)                     # no targets: predictions only
    assert torch.equal(pred3, ref) and int(c3.item()) == 0
