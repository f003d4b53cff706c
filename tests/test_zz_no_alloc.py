"""b200ml_layernorm_bwd_noalloc — the allocation-free LayerNorm backward — against b200ml_layernorm_bwd.

Some users of libb200ml.so run inside a host process that owns all device memory through its own pool and is not
allowed to call cudaMalloc/cudaFree once it is up (the allocation can fail, and it synchronises the GPU), or they call
the C ABI from plain C / an FFI with no allocator to carve a workspace from.  For them the library carries its own
scratch (a static device array).  This test goes through the C ABI only (ctypes on the shared object, raw pointers and
a stream; torch is used just to own the input/output buffers and to compare), calls the new entry point twice in a row
with different shapes — the shared scratch still holds the first call's partial sums during the second — and expects
the same dz / da / dgamma / dbeta as the scratch-taking entry point, within the tolerance the other LayerNorm tests use.

The file name sorts last on purpose: under ``pytest -x`` every other test has already run if this one fails.  It must
pass both alone in a fresh process and after the rest of the suite.
"""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "machine_learning---apache-spark_b200",
                   "lib", "libb200ml.so")


def _load():
    lib = C.CDLL(LIB)
    vp, i, f, u, sz = C.c_void_p, C.c_int, C.c_float, C.c_uint, C.c_size_t
    lib.b200ml_last_error.restype = C.c_char_p
    lib.b200ml_layernorm_bwd_scratch_bytes.argtypes = [i, i]
    lib.b200ml_layernorm_bwd_scratch_bytes.restype = sz
    lib.b200ml_layernorm_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, i, i, f, vp, u, i, vp]
    lib.b200ml_layernorm_bwd.restype = i
    lib.b200ml_layernorm_bwd_noalloc.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, f, vp, u, i, vp]
    lib.b200ml_layernorm_bwd_noalloc.restype = i
    lib.b200ml_layernorm_bwd_noalloc_max_d.restype = sz
    return lib


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _case(lib, rows, d, p, seed):
    """Runs both entry points on identical inputs; returns the two result tuples."""
    torch.manual_seed(seed)
    dev = "cuda"
    dy = torch.randn(rows, d, device=dev).to(torch.bfloat16)
    z = torch.randn(rows, d, device=dev).to(torch.bfloat16)
    mean = z.float().mean(-1).contiguous()
    rstd = torch.rsqrt(z.float().var(-1, unbiased=False) + 1e-5).contiguous()
    gamma = (1 + 0.1 * torch.randn(d, device=dev)).contiguous()
    rng = torch.tensor([1234, 7], dtype=torch.int64, device=dev)   # {seed, step} of the dropout stream
    stream = torch.cuda.current_stream().cuda_stream
    outs = []
    for variant in ("scratch", "noalloc"):
        dz = torch.empty_like(dy)
        da = torch.zeros_like(dy)   # only written when dropout_p > 0 (otherwise d_a is dz itself)
        dg = torch.full((d,), float("nan"), device=dev)
        db = torch.full((d,), float("nan"), device=dev)
        if variant == "scratch":
            n = lib.b200ml_layernorm_bwd_scratch_bytes(rows, d)
            ws = torch.empty(n, dtype=torch.uint8, device=dev)
            rc = lib.b200ml_layernorm_bwd(dy.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                          gamma.data_ptr(), dz.data_ptr(), da.data_ptr(), dg.data_ptr(), db.data_ptr(),
                                          ws.data_ptr(), n, rows, d, p, rng.data_ptr(), 5, 0, stream)
        else:
            rc = lib.b200ml_layernorm_bwd_noalloc(dy.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                                  gamma.data_ptr(), dz.data_ptr(), da.data_ptr(), dg.data_ptr(),
                                                  db.data_ptr(), rows, d, p, rng.data_ptr(), 5, 0, stream)
        assert rc == 0, lib.b200ml_last_error().decode()
        torch.cuda.synchronize()
        outs.append((dz, da, dg, db))
    return outs


def test_noalloc_matches_scratch_variant_on_two_calls_in_a_row():
    lib = _load()
    assert lib.b200ml_layernorm_bwd_noalloc_max_d() == 1024
    # two back-to-back calls with different shapes and dropout settings: the second one finds the first one's partial
    # sums in the shared scratch and must not be affected by them
    for rows, d, p, seed in ((6400, 512, 0.1, 0), (333, 1024, 0.0, 1)):
        (dz0, da0, dg0, db0), (dz1, da1, dg1, db1) = _case(lib, rows, d, p, seed)
        assert _rel(dz1, dz0) < 0.02
        if p > 0:
            assert _rel(da1, da0) < 0.02
        else:
            assert da0.abs().sum().item() == 0 and da1.abs().sum().item() == 0   # untouched by both variants
        assert _rel(dg1, dg0) < 0.02 and _rel(db1, db0) < 0.02
        assert torch.isfinite(dg1).all() and torch.isfinite(db1).all()
        if p > 0:   # the regenerated dropout mask is the same draw in both variants
            assert torch.equal(da1 == 0, da0 == 0)


def test_noalloc_rejects_sizes_beyond_its_scratch():
    lib = _load()
    t = torch.zeros(8, 2048, device="cuda", dtype=torch.bfloat16)
    v = torch.zeros(2048, device="cuda")
    r = torch.zeros(8, device="cuda")
    rc = lib.b200ml_layernorm_bwd_noalloc(t.data_ptr(), t.data_ptr(), r.data_ptr(), r.data_ptr(), v.data_ptr(),
                                          t.data_ptr(), None, v.data_ptr(), v.data_ptr(), 8, 2048, 0.0, None, 0, 0,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.b200ml_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "1024" in msg
