"""Microbenchmark: b200ml tcgen05 GEMM vs torch.matmul (cuBLAS) on the Transformer-MT shapes.
CUDA events, warm-up, L2 flushed between timed iterations.  Writes gpurun_out/bench_gemm.json."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.ops.gemm import gemm  # noqa: E402

SHAPES = [  # (name, M, N, K, a_mn, b_mn)
    ("qkv_fwd", 6400, 1536, 512, False, False),
    ("proj_fwd", 6400, 512, 512, False, False),
    ("ffn1_fwd", 6400, 1024, 512, False, False),
    ("ffn2_fwd", 6400, 512, 1024, False, False),
    ("head_fwd", 6400, 19214, 512, False, False),
    ("head_dgrad", 6400, 512, 19214, False, True),
    ("head_wgrad", 19214, 512, 6400, True, True),
    ("qkv_dgrad", 6400, 512, 1536, False, True),
    ("qkv_wgrad", 1536, 512, 6400, True, True),
    ("square_8k", 8192, 8192, 8192, False, False),
]


def timeit(fn, iters=10, warm=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    res = []
    for name, M, N, K, a_mn, b_mn in SHAPES:
        ld_n = (N + 7) // 8 * 8
        a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(torch.bfloat16)
        if a_mn and M % 8:
            a = torch.randn(K, (M + 7) // 8 * 8, device="cuda").to(torch.bfloat16)[:, :M]
        b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(torch.bfloat16)
        if (not b_mn) and K % 8:
            b = torch.randn(N, (K + 7) // 8 * 8, device="cuda").to(torch.bfloat16)[:, :K]
        if (not a_mn) and K % 8:
            a = torch.randn(M, (K + 7) // 8 * 8, device="cuda").to(torch.bfloat16)[:, :K]
        out = torch.empty(M, ld_n, dtype=torch.bfloat16, device="cuda")
        row = {"name": name, "M": M, "N": N, "K": K}
        for tn in (0, 128, 256):
            try:
                t = timeit(lambda: gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, n_logical=N, tile_n=tn))
                row[f"b200ml_tile{tn}_ms"] = t
                row[f"b200ml_tile{tn}_tflops"] = 2.0 * M * N * K / t / 1e9
            except Exception as ex:  # noqa: BLE001
                row[f"b200ml_tile{tn}_err"] = str(ex)
        A = (a.t() if a_mn else a)
        B = (b.t() if b_mn else b)
        t = timeit(lambda: torch.matmul(A, B.t()))
        row["cublas_ms"] = t
        row["cublas_tflops"] = 2.0 * M * N * K / t / 1e9
        print(json.dumps(row), flush=True)
        res.append(row)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/bench_gemm.json", "w") as fh:
        json.dump(res, fh, indent=1)


if __name__ == "__main__":
    main()
