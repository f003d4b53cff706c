"""Small single-kernel drivers for ncu captures (one GPU, a handful of launches).

    python bench/profile_target.py attn      # attention fwd + bwd at the translator shape
    python bench/profile_target.py gemm_head # 6400 x 19214 x 512 head GEMM with bias
    python bench/profile_target.py gemm_small
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.ops.attention import attention  # noqa: E402
from b200ml.ops.gemm import gemm  # noqa: E402


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "attn"
    torch.manual_seed(0)
    if what == "attn":
        qkv = (torch.randn(32, 200, 1536, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_()
        for _ in range(3):
            out = attention(qkv, num_heads=8, mask_mode="reference")
            out.float().sum().backward()
    elif what == "gemm_head":
        a = torch.randn(6400, 512, device="cuda").to(torch.bfloat16)
        b = torch.randn(19214, 512, device="cuda").to(torch.bfloat16)
        bias = torch.randn(19214, device="cuda")
        out = torch.empty(6400, 19216, dtype=torch.bfloat16, device="cuda")
        for _ in range(3):
            gemm(a, b, out=out, bias=bias, n_logical=19214)
    elif what == "head":
        # the translator's output projection + loss as one autograd node: pair GEMM (fwd), row-buffered loss kernel,
        # pair GEMM (dgrad), pair GEMM (wgrad)
        from b200ml.ops.loss import linear_cross_entropy
        h = (torch.randn(6400, 512, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_()
        w = (torch.randn(19214, 512, device="cuda") * 0.05).requires_grad_()
        b = (torch.randn(19214, device="cuda") * 0.05).requires_grad_()
        t = torch.randint(0, 19214, (6400,), device="cuda")
        for _ in range(2):
            linear_cross_entropy(h, w, b, t, ignore_index=0).backward()
    elif what == "gemm_wgrad":
        # weight-gradient shapes (both operands MN-major, K = B*S = 6400) and a long-K NN shape at the narrow tile
        a = torch.randn(6400, 1536, device="cuda").to(torch.bfloat16)
        b = torch.randn(6400, 512, device="cuda").to(torch.bfloat16)
        out = torch.empty(1536, 512, device="cuda")
        a2 = torch.randn(6400, 4096, device="cuda").to(torch.bfloat16)
        b2 = torch.randn(512, 4096, device="cuda").to(torch.bfloat16)
        for _ in range(2):
            gemm(a, b, a_mn=True, b_mn=True, out=out, tile_n=64, split_k=1)
            gemm(a, b, a_mn=True, b_mn=True, out=out, tile_n=256, split_k=1)
            gemm(a, b, a_mn=True, b_mn=True, out=out, tile_n=128)
            gemm(a2, b2, tile_n=64)
            gemm(a2, b2, tile_n=256)
    else:
        a = torch.randn(6400, 512, device="cuda").to(torch.bfloat16)
        b = torch.randn(1024, 512, device="cuda").to(torch.bfloat16)
        bias = torch.randn(1024, device="cuda")
        for _ in range(3):
            gemm(a, b, bias=bias, act="relu", dropout_p=0.1, site=3)
    torch.cuda.synchronize()
    print("done", what)


if __name__ == "__main__":
    main()
