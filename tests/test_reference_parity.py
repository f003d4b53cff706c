"""CPU oracle tests against the reference classes themselves (transformer.py imported; the other model classes are
ast-extracted because their modules download datasets at import time) — SURVEY §4."""
import ast
import importlib.util
import os

import pytest
import torch


def _extract(reference_dir, path, cls, extra=None):
    src = open(os.path.join(reference_dir, path)).read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.ClassDef) and n.name == cls)
    ns = {"torch": torch, "nn": torch.nn}
    ns.update(extra or {})
    exec(compile(ast.Module([node], []), path, "exec"), ns)
    return ns[cls]


@pytest.fixture(scope="module")
def ref_transformer(reference_dir):
    spec = importlib.util.spec_from_file_location("ref_transformer", os.path.join(reference_dir, "transformer.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class _V:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


def test_transformer_state_dict_and_forward_match(ref_transformer):
    from b200ml.models.transformer import Transformer
    torch.manual_seed(0)
    S, Ve, Vd = 24, 50, 60
    rm = ref_transformer.Transformer(64, 128, 2, 0.1, 2, S, Vd, _V(Ve), _V(Vd)).eval()
    m = Transformer(64, 128, 2, 0.1, 2, S, Vd, _V(Ve), _V(Vd)).eval()
    assert list(rm.state_dict().keys()) == list(m.state_dict().keys())
    assert all(a.shape == b.shape for a, b in zip(rm.state_dict().values(), m.state_dict().values()))
    m.load_state_dict(rm.state_dict(), strict=True)
    rm.load_state_dict(m.state_dict(), strict=True)       # both directions
    x, y = torch.randint(1, Ve, (3, S)), torch.randint(1, Vd, (3, S))
    x[:, -5:] = 0
    pad = (x != 0).unsqueeze(1).unsqueeze(2)
    la = (torch.tril(torch.ones(S, S)) == 0).unsqueeze(0).unsqueeze(0)
    with torch.no_grad():
        o_ref, o = rm(x, y, pad, la, la), m(x, y, pad, la, la)
    assert (o_ref - o).abs().max().item() < 1e-5


def test_attention_reference_mode_equals_reference_sdp(ref_transformer):
    from b200ml.ops.attention import attention_reference
    torch.manual_seed(1)
    B, H, S, hd = 2, 2, 10, 8
    qkv = torch.randn(B, S, 3 * H * hd)
    t = qkv.reshape(B, S, H, 3 * hd).permute(0, 2, 1, 3)
    q, k, v = t.chunk(3, dim=-1)
    la = (torch.tril(torch.ones(S, S)) == 0).unsqueeze(0).unsqueeze(0)
    vals, _ = ref_transformer.scaled_dot_product(q, k, v, la)
    ref = vals.permute(0, 2, 1, 3).reshape(B, S, H * hd)
    assert (attention_reference("self", qkv, None, H, "reference") - ref).abs().max().item() < 1e-5
    pad = torch.ones(B, 1, 1, S, dtype=torch.bool)
    pad[0, ..., -3:] = False
    vals, _ = ref_transformer.scaled_dot_product(q, k, v, pad)      # padding mask is a softmax no-op
    ref = vals.permute(0, 2, 1, 3).reshape(B, S, H * hd)
    assert (attention_reference("self", qkv, None, H, "none") - ref).abs().max().item() < 1e-5


def test_positional_encoding_and_layernorm_match(ref_transformer):
    from b200ml.models.transformer import LayerNormalization, PositionalEncoding
    assert torch.allclose(PositionalEncoding(32, 20)(), ref_transformer.PositionalEncoding(32, 20)(), atol=1e-6)
    torch.manual_seed(0)
    x = torch.randn(4, 7, 32)
    a, b = LayerNormalization([32]), ref_transformer.LayerNormalization([32])
    with torch.no_grad():
        b.gamma.mul_(1.3)
        b.beta.add_(0.2)
    a.load_state_dict(b.state_dict())
    assert torch.allclose(a(x), b(x), atol=1e-5)


def test_cnn_mlp_lstm_state_dicts_and_forward(reference_dir):
    from b200ml.models import LSTM, FashionMNISTModel, Multilayer_perceptor
    torch.manual_seed(0)
    r = _extract(reference_dir, "pytorch_cnn.py", "FashionMNISTModel")(1, 10, 10)
    m = FashionMNISTModel(1, 10, 10)
    assert {k: v.shape for k, v in r.state_dict().items()} == {k: v.shape for k, v in m.state_dict().items()}
    m.load_state_dict(r.state_dict(), strict=True)
    x = torch.randn(4, 1, 28, 28)
    assert torch.allclose(m(x), r(x), atol=1e-5)
    FashionMNISTModel(3, 10, 10, image_size=32)(torch.randn(2, 3, 32, 32))          # CIFAR preset works (ref fails)

    r = _extract(reference_dir, "pytorch_multilayer_perceptron.py", "Multilayer_perceptor")()
    m = Multilayer_perceptor()
    m.load_state_dict(r.state_dict(), strict=True)
    x = torch.randn(30, 4)
    assert torch.allclose(m(x), r(x), atol=1e-6)

    RefLSTM = _extract(reference_dir, "pytorch_lstm.py", "LSTM", {"vocab": {"0": 0}})
    r = RefLSTM(300, 32, 32, 4).eval()
    m = LSTM(300, 32, 32, 4).eval()
    assert list(r.state_dict().keys()) == list(m.state_dict().keys())
    m.load_state_dict(r.state_dict(), strict=True)
    tok = torch.randint(0, 300, (5, 12))
    z = torch.zeros(2, 5, 32)
    (o1, h1, c1), (o2, h2, c2) = r(tok, z, z), m(tok, z, z)
    assert torch.allclose(o1, o2, atol=1e-5) and torch.allclose(h1, h2, atol=1e-5) and torch.allclose(c1, c2, atol=1e-5)


def test_state_dict_torch_save_roundtrip(tmp_path):
    from b200ml.models import Multilayer_perceptor
    from b200ml.ops.optim import FlatOptimizer
    m = Multilayer_perceptor()
    FlatOptimizer(m.parameters(), kind="sgd", lr=0.1)     # parameters become views into the flat buffer
    torch.save(m.state_dict(), tmp_path / "m.pt")
    sd = torch.load(tmp_path / "m.pt")
    assert all(v.dtype == torch.float32 for v in sd.values())
    m2 = Multilayer_perceptor()
    m2.load_state_dict(sd, strict=True)
    assert torch.equal(m2.layer_1.weight, m.layer_1.weight)
