import os, sys, torch
sys.path.insert(0, '/root/repo/bench'); sys.path.insert(0, '/root/repo')
from b200ml.ops.gemm import gemm
from bench_gemm_overhead import timed
a = torch.randn(6400, 512, device="cuda").to(torch.bfloat16)
b = torch.randn(19214, 512, device="cuda").to(torch.bfloat16)
out = torch.empty(6400, 19216, dtype=torch.bfloat16, device="cuda")[:, :19214]
bias = torch.randn(19214, device="cuda")
print("debug", os.environ.get("B200ML_GEMM_DEBUG", "0"), "pair %.1f us" % timed(lambda: gemm(a, b, out=out, tile_n=512)),
      "pair+bias %.1f us" % timed(lambda: gemm(a, b, out=out, bias=bias, tile_n=512)),
      "single %.1f us" % timed(lambda: gemm(a, b, out=out, tile_n=256)))
