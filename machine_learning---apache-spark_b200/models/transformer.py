"""Encoder–decoder Transformer with the reference's class names, constructor signatures and state_dict keys
(reference transformer.py:8-284), executed by the b200ml kernels:

  SentenceEmbedding      -> one fused gather + positional-encoding + dropout kernel (PE table resident on the device)
  MultiHeadAttention     -> qkv GEMM (tcgen05) -> attention kernel reading q/k/v in place from the interleaved layout
  LayerNormalization     -> fused dropout + residual + LayerNorm (params ``gamma`` / ``beta``, biased variance)
  PositionwiseFeedForward-> two tcgen05 GEMMs with bias+ReLU+dropout in the first epilogue, ReLU/dropout derivative
                            folded into the dgrad epilogue on the way back
  Transformer.linear     -> head GEMM; training scripts use ops.linear_cross_entropy on ``return_hidden=True`` output

``mask_mode``: "reference" reproduces the reference as written — boolean masks are *added* to the scores,
so the look-ahead mask is +1.0 on the strictly lower triangle and the padding mask is a no-op; "causal" applies a real
causal mask in decoder self-attention.  Post-LN residual structure, dropout placement and the un-shifted decoder input
are kept exactly.
"""
from __future__ import annotations

import math
import os

import torch
from torch import nn

from ..ops import attention as _att
from ..ops.embedding import embedding as _embedding
from ..ops.embedding import positional_encoding_table
from ..ops.gemm import linear as _linear
from ..ops.gemm import mlp_chain as _mlp_chain
from ..ops.norm import layernorm_residual as _ln_res
from ..utils import workspace as _workspace
from ..utils.rng import new_site


def get_device():
    return torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")


def _vocab_len(v) -> int:
    return int(v) if isinstance(v, int) else len(v)


def scaled_dot_product(q, k, v, mask=None):
    """API-compatible restatement of transformer.py:12-25 for [B,H,S,hd] tensors (generic additive mask).  Returns
    (values, attention).  Unlike the reference, ``mask=None`` works for any batch size."""
    d_k = q.size(-1)
    scaled = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(d_k)
    if mask is not None:
        scaled = scaled + mask.permute(0, 1, 3, 2)
    attention = torch.softmax(scaled.float(), dim=-1).to(q.dtype)
    return torch.matmul(attention, v), attention


def _mask_kind(mask, model_mode: str, is_decoder_self: bool) -> str | torch.Tensor:
    """Map a reference-style mask argument onto a kernel bias mode (shape-based; see module docstring)."""
    if model_mode == "causal":
        return "causal" if is_decoder_self else "none"
    if mask is None:
        return "none"
    if mask.dtype == torch.bool and mask.dim() == 4:
        if mask.shape[-2] == 1:                    # padding mask [B,1,1,S]: per-query constant after the permute
            return "none"
        if mask.shape[0] == 1 and mask.shape[1] == 1 and mask.shape[-1] == mask.shape[-2]:
            return "reference"                     # look-ahead mask [1,1,S,S]
    return mask                                    # anything else: generic additive path


class PositionalEncoding(nn.Module):
    def __init__(self, d_model, max_sequence_length):
        super().__init__()
        self.max_sequence_length = max_sequence_length
        self.d_model = d_model

    def forward(self, device=None):
        return positional_encoding_table(self.max_sequence_length, self.d_model, device or "cpu")


class SentenceEmbedding(nn.Module):
    def __init__(self, max_sequence_length, d_model, input_vocab):
        super().__init__()
        self.vocab_size = _vocab_len(input_vocab)
        self.max_sequence_length = max_sequence_length
        self.embedding = nn.Embedding(self.vocab_size, d_model)
        self.embedding.weight._b200_sparse_rows = True   # row-sparse gradient: the DP engine skips untouched rows
        self.language_to_index = input_vocab
        self.position_encoder = PositionalEncoding(d_model, max_sequence_length)
        self.dropout = nn.Dropout(p=0.1)
        self._site = new_site()

    def forward(self, x):
        w = self.embedding.weight
        assert x.shape[-1] <= self.max_sequence_length, "sequence longer than max_sequence_length"
        pe = self.position_encoder(w.device)
        out_dtype = torch.bfloat16 if w.is_cuda else torch.float32
        return _embedding(x, w, pe=pe, dropout_p=self.dropout.p, site=self._site, out_dtype=out_dtype,
                          training=self.training)


class MultiHeadAttention(nn.Module):
    def __init__(self, d_model, num_heads):
        super().__init__()
        self.d_model = d_model
        self.num_heads = num_heads
        self.head_dim = d_model // num_heads
        self.qkv_layer = nn.Linear(d_model, 3 * d_model)
        self.linear_layer = nn.Linear(d_model, d_model)

    def forward(self, x, mask=None, mask_mode: str = "reference", decoder: bool = False, fork: bool = False,
                skip_out_bias_grad: bool = False):
        """``fork`` / ``skip_out_bias_grad`` are used by the encoder / decoder layers: with ``fork`` the result is
        ``(out, x_skip)`` where ``x_skip`` is ``x`` routed through the qkv projection's autograd node (its backward adds
        the residual branch's gradient in the dgrad GEMM epilogue); ``skip_out_bias_grad`` leaves the output
        projection's bias gradient to the LayerNorm that consumes ``out`` (ops/norm.py ``a_bias``)."""
        kind = _mask_kind(mask, mask_mode, decoder)
        x_skip = None
        if fork:
            qkv, x_skip = _linear(x, self.qkv_layer.weight, self.qkv_layer.bias, fork=True)
        else:
            qkv = _linear(x, self.qkv_layer.weight, self.qkv_layer.bias)
        if isinstance(kind, torch.Tensor):
            values = _generic_attention(qkv, None, self.num_heads, kind)
        else:
            values = _att.attention(qkv, num_heads=self.num_heads, mask_mode=kind)
        out = _linear(values, self.linear_layer.weight, self.linear_layer.bias, skip_bias_grad=skip_out_bias_grad)
        return (out, x_skip) if fork else out


def _generic_attention(qsrc, kvsrc, H, mask):
    kind = "self" if kvsrc is None else "cross"
    d = qsrc.shape[-1] // (3 if kind == "self" else 1)
    q, k, v = _att._split(kind, qsrc, kvsrc, H, d // H)
    values, _ = scaled_dot_product(q, k, v, mask.to(q.dtype) if mask.dtype != torch.bool else mask.to(q.dtype))
    B, _, S, _ = values.shape
    return values.permute(0, 2, 1, 3).reshape(B, S, d)


class LayerNormalization(nn.Module):
    def __init__(self, parameters_shape, eps=1e-5):
        super().__init__()
        self.parameters_shape = parameters_shape
        self.eps = eps
        self.gamma = nn.Parameter(torch.ones(parameters_shape))
        self.beta = nn.Parameter(torch.zeros(parameters_shape))

    @property
    def fusable(self) -> bool:
        """True when the fused kernel path is taken (1-D parameters): only then may a producer leave its bias
        gradient to this LayerNorm (``a_bias``)."""
        return len(self.parameters_shape) == 1

    def forward(self, inputs, residual=None, dropout_p: float = 0.0, site: int = 0, a_bias=None):
        """LN(dropout(inputs) + residual); with the defaults this is the reference's plain LayerNormalization."""
        if len(self.parameters_shape) != 1:
            assert a_bias is None
            dims = [-(i + 1) for i in range(len(self.parameters_shape))]
            z = inputs if residual is None else inputs + residual
            mean = z.mean(dim=dims, keepdim=True)
            var = ((z - mean) ** 2).mean(dim=dims, keepdim=True)
            return self.gamma * ((z - mean) / (var + self.eps).sqrt()) + self.beta
        return _ln_res(inputs, residual, self.gamma, self.beta, self.eps, dropout_p, site, self.training, a_bias)


class PositionwiseFeedForward(nn.Module):
    def __init__(self, d_model, hidden, drop_prob=0.1):
        super().__init__()
        self.linear1 = nn.Linear(d_model, hidden)
        self.linear2 = nn.Linear(hidden, d_model)
        self.relu = nn.ReLU()
        self.dropout = nn.Dropout(p=drop_prob)
        self._site = new_site()

    def forward(self, x, fork: bool = False, skip_out_bias_grad: bool = False):
        """With ``fork`` the result is ``(out, x_skip)``; see MultiHeadAttention.forward."""
        return _mlp_chain(x, [self.linear1.weight, self.linear2.weight], [self.linear1.bias, self.linear2.bias],
                          ["relu", None], [self.dropout.p, 0.0], [self._site, 0], self.training, fork=fork,
                          skip_last_bias_grad=skip_out_bias_grad)


class EncoderLayer(nn.Module):
    def __init__(self, d_model, ffn_hidden, num_heads, drop_prob):
        super().__init__()
        self.attention = MultiHeadAttention(d_model=d_model, num_heads=num_heads)
        self.norm1 = LayerNormalization(parameters_shape=[d_model])
        self.dropout1 = nn.Dropout(p=drop_prob)
        self.ffn = PositionwiseFeedForward(d_model=d_model, hidden=ffn_hidden, drop_prob=drop_prob)
        self.norm2 = LayerNormalization(parameters_shape=[d_model])
        self.dropout2 = nn.Dropout(p=drop_prob)
        self._sites = (new_site(), new_site())
        self.mask_mode = "reference"

    def forward(self, x, self_attention_mask=None):
        # Each sub-layer is LN(dropout(sub(x)) + x).  `x` reaches the LayerNorm through the sub-layer's first GEMM node
        # (fork), so backward adds the two gradients of `x` in that GEMM's epilogue instead of in an elementwise
        # kernel, and the bias gradient of the sub-layer's last Linear comes out of the LayerNorm backward kernel.
        f1, f2 = self.norm1.fusable, self.norm2.fusable
        a, x_skip = self.attention(x, mask=self_attention_mask, mask_mode=self.mask_mode, decoder=False, fork=True,
                                   skip_out_bias_grad=f1)
        x = self.norm1(a, x_skip, self.dropout1.p, self._sites[0],
                       a_bias=self.attention.linear_layer.bias if f1 else None)
        f, x_skip = self.ffn(x, fork=True, skip_out_bias_grad=f2)
        return self.norm2(f, x_skip, self.dropout2.p, self._sites[1], a_bias=self.ffn.linear2.bias if f2 else None)


class SequentialEncoder(nn.Sequential):
    def forward(self, *inputs):
        x, self_attention_mask = inputs
        for module in self._modules.values():
            x = module(x, self_attention_mask)
        return x


class Encoder(nn.Module):
    def __init__(self, d_model, ffn_hidden, num_heads, drop_prob, num_layers: int, max_sequence_length, input_vocab):
        super().__init__()
        self.sentence_embedding = SentenceEmbedding(max_sequence_length, d_model, input_vocab)
        self.layers = SequentialEncoder(*[EncoderLayer(d_model, ffn_hidden, num_heads, drop_prob)
                                          for _ in range(num_layers)])

    def forward(self, x, self_attention_mask=None):
        x = self.sentence_embedding(x)
        return self.layers(x, self_attention_mask)


class MultiHeadCrossAttention(nn.Module):
    def __init__(self, d_model, num_heads):
        super().__init__()
        self.d_model = d_model
        self.num_heads = num_heads
        self.head_dim = d_model // num_heads
        self.kv_layer = nn.Linear(d_model, 2 * d_model)
        self.q_layer = nn.Linear(d_model, d_model)
        self.linear_layer = nn.Linear(d_model, d_model)

    def forward(self, x, y, mask=None, mask_mode: str = "reference", fork: bool = False,
                skip_out_bias_grad: bool = False):
        """x: encoder output (keys/values), y: decoder stream (queries).  ``fork``: returns ``(out, y_skip)`` with
        ``y`` routed through the query projection's autograd node; see MultiHeadAttention.forward."""
        kind = _mask_kind(mask, mask_mode, False)
        kv = _linear(x, self.kv_layer.weight, self.kv_layer.bias)
        y_skip = None
        if fork:
            q, y_skip = _linear(y, self.q_layer.weight, self.q_layer.bias, fork=True)
        else:
            q = _linear(y, self.q_layer.weight, self.q_layer.bias)
        if isinstance(kind, torch.Tensor):
            values = _generic_attention(q, kv, self.num_heads, kind)
        else:
            values = _att.attention(q, kv, num_heads=self.num_heads, mask_mode=kind)
        out = _linear(values, self.linear_layer.weight, self.linear_layer.bias, skip_bias_grad=skip_out_bias_grad)
        return (out, y_skip) if fork else out


class DecoderLayer(nn.Module):
    def __init__(self, d_model, ffn_hidden, num_heads, drop_prob):
        super().__init__()
        self.self_attention = MultiHeadAttention(d_model=d_model, num_heads=num_heads)
        self.layer_norm1 = LayerNormalization(parameters_shape=[d_model])
        self.dropout1 = nn.Dropout(p=drop_prob)
        self.encoder_decoder_attention = MultiHeadCrossAttention(d_model=d_model, num_heads=num_heads)
        self.layer_norm2 = LayerNormalization(parameters_shape=[d_model])
        self.dropout2 = nn.Dropout(p=drop_prob)
        self.ffn = PositionwiseFeedForward(d_model=d_model, hidden=ffn_hidden, drop_prob=drop_prob)
        self.layer_norm3 = LayerNormalization(parameters_shape=[d_model])
        self.dropout3 = nn.Dropout(p=drop_prob)
        self._sites = (new_site(), new_site(), new_site())
        self.mask_mode = "reference"

    def forward_self(self, y, self_attention_mask=None):
        """Self-attention sub-layer: the part of the decoder layer that does not depend on the encoder output."""
        f1 = self.layer_norm1.fusable
        a, y_skip = self.self_attention(y, mask=self_attention_mask, mask_mode=self.mask_mode, decoder=True, fork=True,
                                        skip_out_bias_grad=f1)
        return self.layer_norm1(a, y_skip, self.dropout1.p, self._sites[0],
                                a_bias=self.self_attention.linear_layer.bias if f1 else None)

    def forward(self, x, y, self_attention_mask=None, cross_attention_mask=None, self_done: bool = False):
        if not self_done:
            y = self.forward_self(y, self_attention_mask)
        f2, f3 = self.layer_norm2.fusable, self.layer_norm3.fusable
        c, y_skip = self.encoder_decoder_attention(x, y, mask=cross_attention_mask, mask_mode=self.mask_mode, fork=True,
                                                   skip_out_bias_grad=f2)
        y = self.layer_norm2(c, y_skip, self.dropout2.p, self._sites[1],
                             a_bias=self.encoder_decoder_attention.linear_layer.bias if f2 else None)
        f, y_skip = self.ffn(y, fork=True, skip_out_bias_grad=f3)
        return self.layer_norm3(f, y_skip, self.dropout3.p, self._sites[2],
                                a_bias=self.ffn.linear2.bias if f3 else None)


class SequentialDecoder(nn.Sequential):
    def forward(self, *inputs, first_self_done: bool = False):
        x, y, self_attention_mask, cross_attention_mask = inputs
        for i, module in enumerate(self._modules.values()):
            y = module(x, y, self_attention_mask, cross_attention_mask, self_done=(first_self_done and i == 0))
        return y


class Decoder(nn.Module):
    def __init__(self, d_model, ffn_hidden, num_heads, drop_prob, num_layers: int, max_sequence_length, input_vocab):
        super().__init__()
        self.sentence_embedding = SentenceEmbedding(max_sequence_length, d_model, input_vocab)
        self.layers = SequentialDecoder(*[DecoderLayer(d_model, ffn_hidden, num_heads, drop_prob)
                                          for _ in range(num_layers)])

    def forward_pre(self, y, self_attention_mask=None):
        """Everything of the decoder that can run before (= concurrently with) the encoder: target embedding and the
        first layer's self-attention sub-layer."""
        y = self.sentence_embedding(y)
        return self.layers[0].forward_self(y, self_attention_mask)

    def forward(self, x, y, self_attention_mask=None, cross_attention_mask=None, pre_done: bool = False):
        if not pre_done:
            y = self.sentence_embedding(y)
        return self.layers(x, y, self_attention_mask, cross_attention_mask, first_self_done=pre_done)


class Transformer(nn.Module):
    def __init__(self, d_model, ffn_hidden, num_heads, drop_prob, num_layers: int, max_sequence_length, de_vocab_size,
                 input_en_vocab, input_dec_vocab, mask_mode: str = "reference"):
        super().__init__()
        self.encoder = Encoder(d_model, ffn_hidden, num_heads, drop_prob, num_layers, max_sequence_length,
                               input_en_vocab)
        self.decoder = Decoder(d_model, ffn_hidden, num_heads, drop_prob, num_layers, max_sequence_length,
                               input_dec_vocab)
        self.linear = nn.Linear(d_model, de_vocab_size)
        self.device = get_device()
        self.set_mask_mode(mask_mode)
        # run the decoder's encoder-independent prefix on a second stream (forward, and therefore backward too)
        self.overlap_streams = os.environ.get("B200ML_OVERLAP_STREAMS", "1") != "0"
        self._dec_stream = None   # lazily created side stream for the decoder prefix (runtime state, never pickled)

    def set_mask_mode(self, mode: str) -> None:
        assert mode in ("reference", "causal")
        self.mask_mode = mode
        for m in self.modules():
            if isinstance(m, (EncoderLayer, DecoderLayer)):
                m.mask_mode = mode


    def __getstate__(self):
        # a returned / pickled model (TorchDistributor ships rank 0's return value to the driver) must not drag CUDA
        # stream handles along
        state = dict(self.__dict__)
        state["_dec_stream"] = None
        return state
    def forward(self, x, y, encoder_self_attention_mask=None, decoder_self_attention_mask=None,
                decoder_cross_attention_mask=None, return_hidden: bool = False):
        if self.overlap_streams and x.is_cuda and len(self.decoder.layers) > 0:
            if self._dec_stream is None:
                self._dec_stream = torch.cuda.Stream(device=x.device)
                _workspace.register_stream(self._dec_stream, 2)
            cur = torch.cuda.current_stream()
            self._dec_stream.wait_stream(cur)       # fork here: the prefix depends on nothing the encoder does
            # Encoder and decoder prefix run side by side on two streams either way; which one is ISSUED first decides
            # the order of backward: autograd runs ready nodes in reverse creation order.
            #  * data-parallel (gradients all-reduced): encoder first.  Backward then issues the decoder prefix (target
            #    embedding last of it) BEFORE the encoder, so the decoder table's bucket - the biggest late one - is
            #    handed to the all-reduce while the encoder's backward is still running, and only the encoder's bucket
            #    is left for the end of the step (the other way round it was queued behind the encoder table's and
            #    both sat in the exposed tail).  The all-reduce is NVLink-bound on few SMs: the overlap is nearly free.
            #  * one GPU: decoder prefix first.  The per-bucket optimizer kernel is HBM-bound on every SM; run under
            #    the encoder's backward it slows that down by more than it hides (measured 1.464 vs 1.434 ms/step).
            enc_first = os.environ.get("B200ML_ENC_FIRST", "auto")
            enc_first = getattr(self, "_b200_dp_world", 1) > 1 if enc_first == "auto" else enc_first == "1"
            if enc_first:
                x = self.encoder(x, encoder_self_attention_mask)
            with torch.cuda.stream(self._dec_stream):
                y_pre = self.decoder.forward_pre(y, decoder_self_attention_mask)
            if not enc_first:
                x = self.encoder(x, encoder_self_attention_mask)
            cur.wait_stream(self._dec_stream)
            y_pre.record_stream(cur)
            out = self.decoder(x, y_pre, decoder_self_attention_mask, decoder_cross_attention_mask, pre_done=True)
        else:
            x = self.encoder(x, encoder_self_attention_mask)
            out = self.decoder(x, y, decoder_self_attention_mask, decoder_cross_attention_mask)
        if return_hidden:
            return out
        return _linear(out, self.linear.weight, self.linear.bias)
