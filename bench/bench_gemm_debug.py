"""Timing experiment: per-k-block cost of the GEMM pipeline with the TMA loads and/or the MMAs switched off
(B200ML_GEMM_DEBUG = 1 / 2 / 3; results are garbage, only the time matters)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.ops.gemm import gemm
from bench_gemm_overhead import timed
for (M, N, K, a_mn, b_mn, tn) in [(6400, 512, 4096, False, False, 64), (6400, 512, 4096, False, False, 128), (6400, 512, 4096, False, False, 256),
                                  (6400, 512, 4096, False, False, 512), (1536, 512, 6400, True, True, 64), (512, 512, 6400, True, True, 64)]:
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(torch.bfloat16)
    out = torch.empty(M, N, dtype=torch.float32, device="cuda")
    t = timed(lambda: gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, tile_n=tn, split_k=1))
    print(f"debug={os.environ.get('B200ML_GEMM_DEBUG','0')} M={M} N={N} K={K} {'T' if a_mn else 'N'}{'T' if b_mn else 'N'} bn{tn}: {t:7.2f} us  ({t/((K+63)//64):.3f} us per k-block incl. fixed)", flush=True)
