"""Ops: thin autograd wrappers around the sm_100a kernels, each with a PyTorch restatement used on CPU tensors
(and as the numerics oracle in tests).  CUDA tensors always take the native path."""
from .gemm import gemm, linear, mlp_chain  # noqa: F401
from .norm import layernorm_residual  # noqa: F401
from .embedding import embedding, positional_encoding_table  # noqa: F401
from .loss import cross_entropy, linear_cross_entropy  # noqa: F401
