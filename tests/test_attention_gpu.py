"""tcgen05 attention kernels vs the fp32 PyTorch restatement (which itself is checked against the reference's
scaled_dot_product in tests/test_reference_parity.py)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("mask_mode", ["none", "reference", "causal"])
@pytest.mark.parametrize("B,H,S", [(2, 2, 64), (2, 8, 200), (1, 3, 129), (3, 1, 256)])
def test_self_attention(mask_mode, B, H, S):
    from b200ml.ops.attention import attention, attention_reference
    torch.manual_seed(0)
    d = H * 64
    qkv = (torch.randn(B, S, 3 * d, device="cuda") * 0.7).to(torch.bfloat16).requires_grad_()
    out = attention(qkv, num_heads=H, mask_mode=mask_mode)
    w = torch.randn(B, S, d, device="cuda")
    (out.float() * w).sum().backward()
    ref_in = qkv.detach().float().cpu().requires_grad_()
    ref = attention_reference("self", ref_in, None, H, mask_mode)
    (ref * w.cpu().to(torch.bfloat16).float()).sum().backward()
    assert _rel(out, ref) < 0.02, _rel(out, ref)
    assert _rel(qkv.grad, ref_in.grad) < 0.04, _rel(qkv.grad, ref_in.grad)


@pytest.mark.parametrize("mask_mode", ["reference", "none"])
@pytest.mark.parametrize("Sq,Skv", [(200, 200), (72, 136), (200, 40), (130, 256)])
def test_cross_attention(mask_mode, Sq, Skv):
    from b200ml.ops.attention import attention, attention_reference
    torch.manual_seed(1)
    B, H = 2, 4
    d = H * 64
    q = (torch.randn(B, Sq, d, device="cuda") * 0.7).to(torch.bfloat16).requires_grad_()
    kv = (torch.randn(B, Skv, 2 * d, device="cuda") * 0.7).to(torch.bfloat16).requires_grad_()
    out = attention(q, kv, num_heads=H, mask_mode=mask_mode)
    w = torch.randn(B, Sq, d, device="cuda")
    (out.float() * w).sum().backward()
    qr = q.detach().float().cpu().requires_grad_()
    kvr = kv.detach().float().cpu().requires_grad_()
    ref = attention_reference("cross", qr, kvr, H, mask_mode)
    (ref * w.cpu().to(torch.bfloat16).float()).sum().backward()
    assert _rel(out, ref) < 0.02
    assert _rel(q.grad, qr.grad) < 0.04
    assert _rel(kv.grad, kvr.grad) < 0.04
