"""Fixed cost vs per-k-block slope of the tcgen05 GEMM (diagnostic): R back-to-back launches inside one CUDA graph, so the
host launch path is out of the measurement.  Prints us per GEMM for b200ml (tile_n sweep) and torch.matmul (cuBLAS)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.ops.gemm import gemm  # noqa: E402

R = 40


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(R):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / R)
    return best


def main():
    torch.manual_seed(0)
    shapes = [(6400, 512, K, False, False) for K in (64, 256, 512, 1024, 2048, 4096)]
    shapes += [(6400, 19214, 512, False, False), (6400, 512, 19214, False, True), (19214, 512, 6400, True, True),
               (8192, 8192, 8192, False, False)]
    shapes += [(512, 512, 6400, True, True), (1536, 512, 6400, True, True), (1024, 512, 6400, True, True),
               (6400, 1536, 512, False, False), (6400, 512, 1536, False, True)]
    for M, N, K, a_mn, b_mn in shapes:
        r8 = lambda v: (v + 7) // 8 * 8  # TMA needs 16-byte row pitches: pad the leading dimension, view the logical part
        a = torch.randn((K, r8(M)) if a_mn else (M, r8(K)), device="cuda").to(torch.bfloat16)
        a = a[:, :M] if a_mn else a[:, :K]
        b = torch.randn((K, r8(N)) if b_mn else (N, r8(K)), device="cuda").to(torch.bfloat16)
        b = b[:, :N] if b_mn else b[:, :K]
        ldc = (N + 7) // 8 * 8
        out = torch.empty(M, ldc, dtype=torch.float32 if a_mn else torch.bfloat16, device="cuda")[:, :N]
        row = f"M={M:5d} N={N:5d} K={K:5d} {'T' if a_mn else 'N'}{'T' if b_mn else 'N'}"
        for tn in (0, 128, 256, 512):
            t = timed(lambda: gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, tile_n=tn))
            row += f"  bn{tn}: {t:6.2f}us"
        if a_mn:
            t1 = timed(lambda: gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, split_k=1))
            row += f"  nosplit: {t1:6.2f}us"
        am = a.t() if a_mn else a
        bm = b if b_mn else b.t()
        o2 = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        t = timed(lambda: torch.matmul(am, bm, out=o2))
        row += f"  cublas: {t:6.2f}us"
        print(row, flush=True)


if __name__ == "__main__":
    main()
