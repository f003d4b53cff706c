"""Cost of the fused epilogues at the translator's Linear shapes (diagnostic; CUDA-graph chains of 40 launches)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.ops.gemm import gemm
from bench_gemm_overhead import timed
torch.manual_seed(0)
M = 6400
for N, K in [(512, 512), (1024, 512), (1536, 512), (512, 1024)]:
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    b = torch.randn(N, K, device="cuda").to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    aux = torch.randn(M, N, device="cuda").to(torch.bfloat16)
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    out32 = torch.empty(M, N, dtype=torch.float32, device="cuda")
    rows = {
        "plain bf16": lambda: gemm(a, b, out=out),
        "bias": lambda: gemm(a, b, out=out, bias=bias),
        "bias+relu": lambda: gemm(a, b, out=out, bias=bias, act="relu"),
        "bias+relu+drop": lambda: gemm(a, b, out=out, bias=bias, act="relu", dropout_p=0.1, site=3),
        "bias+drop": lambda: gemm(a, b, out=out, bias=bias, dropout_p=0.1, site=3),
        "aux relu-mask": lambda: gemm(a, b, out=out, aux=aux, aux_mode=1, aux_scale=1.0),
        "fp32 out": lambda: gemm(a, b, out=out32),
        "direct store": lambda: gemm(a, b, out=out, epilogue_direct=True),
    }
    line = f"M={M} N={N} K={K}:"
    for name, fn in rows.items():
        try:
            line += f"  {name} {timed(fn):.2f}"
        except Exception as e:
            line += f"  {name} ERR({str(e)[:40]})"
    print(line, flush=True)
