"""tcgen05 GEMM vs a plain fp32 PyTorch restatement (operands rounded to bf16 first)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    A = (a.t() if a_mn else a).float().cpu()
    B = (b.t() if b_mn else b).float().cpu()
    return A @ B.t()


def _check(out, ref, K, tol_scale=1.0):
    """The operands are exact bf16 values and accumulation is fp32, so an fp32 output must agree with the fp32 oracle
    to rounding level: relative Frobenius error <= 2e-3 and no element further off than 2e-3 of the largest - a dropped
    k-block (1 % of K = 6400), a wrong split-K seam or a misplaced tile all exceed that by far.  A bf16 output adds its
    own rounding (2^-9 relative per element): 6e-3."""
    fp32_out = out.dtype == torch.float32
    out = out.float().cpu()
    tol = (2e-3 if fp32_out else 6e-3) * tol_scale
    scale = ref.abs().max().item() + 1e-12
    rel = ((out - ref).norm() / (ref.norm() + 1e-12)).item()
    worst = (out - ref).abs().max().item() / scale
    assert rel <= tol and worst <= tol, f"relative error {rel:.3e} (norm) / {worst:.3e} (worst element) > {tol:.1e}"


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K,tile_n", [
    (128, 64, 64, 64), (256, 128, 128, 128), (384, 512, 192, 256), (200, 136, 72, 0), (1000, 264, 520, 0),
    (6400, 512, 512, 0),
    # tile_n = 512: CTA-pair (cta_group::2) 256 x 256 tiles, incl. ragged M / N / K and more tiles than pairs
    (256, 256, 64, 512), (512, 512, 192, 512), (1000, 264, 520, 512), (5000, 2048, 136, 512), (200, 136, 72, 512),
])
def test_gemm_layouts(a_mn, b_mn, M, N, K, tile_n):
    from b200ml.ops.gemm import gemm
    torch.manual_seed(0)
    dev = "cuda"
    a = torch.randn((K, M) if a_mn else (M, K), device=dev).to(torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device=dev).to(torch.bfloat16)
    out = gemm(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32, tile_n=tile_n)
    torch.cuda.synchronize()
    _check(out, _ref(a, b, a_mn, b_mn), K)
    out16 = gemm(a, b, a_mn=a_mn, b_mn=b_mn, tile_n=tile_n)
    _check(out16, _ref(a, b, a_mn, b_mn), K, tol_scale=2.0)


def test_gemm_simt_matches():
    from b200ml.ops.gemm import gemm
    torch.manual_seed(1)
    a = torch.randn(77, 50, device="cuda").to(torch.bfloat16)   # lda = 50: not TMA-describable -> SIMT
    b = torch.randn(33, 50, device="cuda").to(torch.bfloat16)
    out = gemm(a, b, out_dtype=torch.float32)
    _check(out, _ref(a, b, False, False), 50)
    a2 = torch.randn(256, 128, device="cuda").to(torch.bfloat16)
    b2 = torch.randn(128, 128, device="cuda").to(torch.bfloat16)
    o1 = gemm(a2, b2, out_dtype=torch.float32, force_simt=True)
    o2 = gemm(a2, b2, out_dtype=torch.float32)
    assert (o1 - o2).abs().max().item() < 1e-2


def test_gemm_simt_split_k_weight_gradient_shape():
    """dW[4, 32] = dZᵀ[4, 4128]·X[4128, 32]: lda = 4 is not TMA-describable and K is long -> SIMT with split-K."""
    from b200ml.ops.gemm import gemm
    torch.manual_seed(11)
    K, M, N = 4128 + 5, 4, 32
    a = torch.randn(K, M, device="cuda").to(torch.bfloat16)
    b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
    ref = _ref(a, b, True, True)
    out = gemm(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32)
    _check(out, ref, K)
    acc = torch.ones(M, N, device="cuda")
    gemm(a, b, a_mn=True, b_mn=True, out=acc, accumulate=True, alpha=0.5)
    _check(acc, 1.0 + 0.5 * ref, K)
    one = gemm(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32, split_k=1)
    _check(one, ref, K)


def test_gemm_pair_epilogue_and_split_k():
    """CTA-pair path with the fused epilogue (bias + ReLU, bf16 TMA store), fp32 accumulate, and split-K."""
    from b200ml.ops.gemm import gemm
    torch.manual_seed(5)
    M, N, K = 1300, 1000, 256
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    b = (torch.randn(N, K, device="cuda") * 0.1).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    ref = torch.relu(_ref(a, b, False, False) + bias.cpu())
    _check(gemm(a, b, bias=bias, act="relu", tile_n=512), ref, K, tol_scale=2.0)
    _check(gemm(a, b, bias=bias, act="relu", tile_n=512, out_dtype=torch.float32), ref, K)
    # long K, few tiles -> split-K over pairs
    K2 = 6400
    a2 = torch.randn(K2, 512, device="cuda").to(torch.bfloat16)
    b2 = torch.randn(K2, 512, device="cuda").to(torch.bfloat16)
    ref2 = _ref(a2, b2, True, True)
    acc = torch.ones(512, 512, device="cuda")
    gemm(a2, b2, a_mn=True, b_mn=True, out=acc, accumulate=True, tile_n=512)
    _check(acc, 1.0 + ref2, K2)


@pytest.mark.parametrize("M,N,K,tile_n", [(512, 512, 6400, 128), (512, 512, 6400, 256), (300, 200, 3000, 0),
                                          (1024, 136, 4096, 64)])
def test_gemm_cluster_split_k(M, N, K, tile_n):
    """Long K, few tiles: split-K inside a thread-block cluster (DSMEM reduction), incl. ragged M / N, fp32
    accumulate with alpha, and a bf16 output."""
    from b200ml.ops.gemm import gemm
    torch.manual_seed(13)
    a = torch.randn(K, M, device="cuda").to(torch.bfloat16)
    b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
    ref = _ref(a, b, True, True)
    out = gemm(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32, tile_n=tile_n)
    _check(out, ref, K)
    one = gemm(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32, tile_n=tile_n, split_k=1)
    _check(one, ref, K)
    acc = torch.full((M, N), 2.0, device="cuda")
    gemm(a, b, a_mn=True, b_mn=True, out=acc, accumulate=True, alpha=0.25, tile_n=tile_n)
    _check(acc, 2.0 + 0.25 * ref, K)
    ldc = (N + 7) // 8 * 8
    o16 = torch.empty(M, ldc, dtype=torch.bfloat16, device="cuda")[:, :N]
    gemm(a, b, a_mn=True, b_mn=True, out=o16, tile_n=tile_n)
    _check(o16, ref, K, tol_scale=2.0)


def test_gemm_pair_tail_split():
    """77 pair tiles on 74 pairs: the first 74 run as full tiles, the last 3 (768 rows, the final one ragged) as a
    split-K GEMM over the idle pairs; the seam must be invisible."""
    from b200ml.ops.gemm import gemm
    torch.manual_seed(9)
    M, N, K = 76 * 256 + 100, 256, 2048
    a = torch.randn(K, M, device="cuda").to(torch.bfloat16)
    b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
    ref = _ref(a, b, True, True)
    out = gemm(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32, tile_n=512)
    _check(out, ref, K)
    a2 = a.t().contiguous()      # K-major A: the tail's row offset goes through lda
    out2 = gemm(a2, b, b_mn=True, out_dtype=torch.float32, tile_n=512)
    _check(out2, ref, K)


def test_gemm_epilogues():
    from b200ml.ops.gemm import gemm
    torch.manual_seed(2)
    M, N, K = 300, 200, 128
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    b = (torch.randn(N, K, device="cuda") * 0.1).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    aux = torch.randn(M, N, device="cuda").to(torch.bfloat16)
    base = a.float().cpu() @ b.float().cpu().t()
    o = gemm(a, b, bias=bias, act="relu", out_dtype=torch.float32)
    _check(o, torch.relu(base + bias.cpu()), K)
    o = gemm(a, b, bias=bias, act="sigmoid", out_dtype=torch.float32)
    _check(o, torch.sigmoid(base + bias.cpu()), K)
    o = gemm(a, b, aux=aux, aux_mode=1, aux_scale=1.25, out_dtype=torch.float32)
    _check(o, torch.where(aux.float().cpu() > 0, base * 1.25, torch.zeros_like(base)), K)
    o = gemm(a, b, aux=aux, aux_mode=2, out_dtype=torch.float32, alpha=0.5)
    _check(o, 0.5 * base + aux.float().cpu(), K)
    acc = torch.ones(M, N, device="cuda")
    gemm(a, b, out=acc, accumulate=True, alpha_dev=torch.tensor([2.0], device="cuda"))
    _check(acc, 2 * base + 1, K)
    # padded output / logical N
    outp = torch.zeros(M, 208, dtype=torch.bfloat16, device="cuda")
    gemm(a, b, out=outp, n_logical=N)
    _check(outp[:, :N], base, K, tol_scale=2.0)
    assert outp[:, N:].abs().max().item() == 0


def test_gemm_dropout_statistics_and_replay():
    from b200ml.ops.gemm import gemm
    from b200ml.utils.rng import DeviceCounters
    ctr = DeviceCounters.get("cuda")
    ctr.seed(123)
    a = torch.ones(512, 64, device="cuda").to(torch.bfloat16)
    b = torch.ones(256, 64, device="cuda").to(torch.bfloat16)
    o1 = gemm(a, b, dropout_p=0.25, site=7, out_dtype=torch.float32)
    o2 = gemm(a, b, dropout_p=0.25, site=7, out_dtype=torch.float32)
    assert torch.equal(o1, o2)                      # same (seed, step, site) -> same mask
    keep = (o1 != 0).float().mean().item()
    assert abs(keep - 0.75) < 0.01
    assert abs(o1.max().item() - 64 / 0.75) < 0.5
    ctr.advance()
    o3 = gemm(a, b, dropout_p=0.25, site=7, out_dtype=torch.float32)
    assert not torch.equal(o1, o3)


def test_linear_chain_grads_match_torch():
    from b200ml.ops.gemm import mlp_chain
    torch.manual_seed(3)
    x = (torch.randn(4, 50, 64, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_()
    w1 = (torch.randn(128, 64, device="cuda") * 0.1).requires_grad_()
    b1 = torch.randn(128, device="cuda").mul_(0.1).requires_grad_()
    w2 = (torch.randn(72, 128, device="cuda") * 0.1).requires_grad_()
    b2 = torch.randn(72, device="cuda").mul_(0.1).requires_grad_()
    for act in ("relu", "sigmoid"):
        for t in (x, w1, b1, w2, b2):
            t.grad = None
        y = mlp_chain(x, [w1, w2], [b1, b2], [act, None])
        y.float().square().sum().backward()
        got = [t.grad.float().cpu().clone() for t in (x, w1, b1, w2, b2)]
        xr, w1r, b1r, w2r, b2r = [t.detach().float().cpu().requires_grad_() for t in (x, w1, b1, w2, b2)]
        w1q, w2q = w1r.to(torch.bfloat16).float(), w2r.to(torch.bfloat16).float()
        # straight-through the bf16 rounding of weights
        h = (torch.relu if act == "relu" else torch.sigmoid)(xr @ (w1r + (w1q - w1r).detach()).t() + b1r)
        yr = h @ (w2r + (w2q - w2r).detach()).t() + b2r
        yr.square().sum().backward()
        for g, r in zip(got, (xr, w1r, b1r, w2r, b2r)):
            rel = (g - r.grad).norm() / (r.grad.norm() + 1e-6)
            assert rel < 0.03, f"{act}: rel grad err {rel}"


@pytest.mark.parametrize("split_k", [0, 4, 7])
def test_gemm_split_k_wgrad_shape(split_k):
    """wgrad of a small Linear: few output tiles, long reduction -> split-K slabs + deterministic reduce."""
    from b200ml.ops.gemm import gemm
    torch.manual_seed(4)
    rows, N, K = 6400, 512, 512
    dy = (torch.randn(rows, N, device="cuda") * 0.1).to(torch.bfloat16)
    x = torch.randn(rows, K, device="cuda").to(torch.bfloat16)
    ref = dy.float().cpu().t() @ x.float().cpu()
    out = gemm(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32, split_k=split_k)
    _check(out, ref, rows)
    out2 = gemm(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32, split_k=split_k)
    assert torch.equal(out, out2)                                  # deterministic
    acc = torch.ones(N, K, device="cuda")
    gemm(dy, x, a_mn=True, b_mn=True, out=acc, accumulate=True, split_k=split_k, alpha=0.5)
    _check(acc, 0.5 * ref + 1, rows)
    o16 = gemm(dy, x, a_mn=True, b_mn=True, split_k=split_k)
    _check(o16, ref, rows, tol_scale=2.0)


@pytest.mark.parametrize("M,N,K", [(6400, 1536, 512), (200, 19214, 512), (333, 200, 128), (128, 136, 64)])
def test_gemm_tma_store_epilogue_matches_direct(M, N, K):
    """bf16 outputs leave through swizzled smem + cp.async.bulk.tensor stores; the register->global path is the oracle
    (edges in M and N, padded leading dimension, bias + ReLU + residual epilogue)."""
    from b200ml.ops.gemm import gemm
    torch.manual_seed(5)
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    b = (torch.randn(N, K, device="cuda") * 0.2).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    ld = (N + 7) // 8 * 8 + 8
    aux = torch.randn(M, ld, device="cuda").to(torch.bfloat16)
    outs = []
    for direct in (False, True):
        out = torch.full((M, ld), 7.0, dtype=torch.bfloat16, device="cuda")
        gemm(a, b, out=out, bias=bias, act="relu", aux=aux, aux_mode=2, n_logical=N, epilogue_direct=direct)
        outs.append(out)
    torch.cuda.synchronize()
    assert torch.equal(outs[0][:, :N], outs[1][:, :N])
    # TMA clips stores at 16-byte granularity: up to 7 padding elements right after column N-1 may be rewritten (always
    # inside the row's own padding, since ldc % 8 == 0); everything beyond stays untouched
    n8 = (N + 7) // 8 * 8
    assert (outs[0][:, n8:] == 7.0).all() and (outs[1][:, N:] == 7.0).all()
    ref = torch.relu(a.float() @ b.float().t() + bias) + aux[:, :N].float()
    _check(outs[0][:, :N], ref.cpu(), K, tol_scale=2.0)
