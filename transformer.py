"""Encoder–decoder Transformer library — same module name and public symbols as the reference's ``transformer.py``
(get_device, scaled_dot_product, PositionalEncoding, SentenceEmbedding, MultiHeadAttention, LayerNormalization,
PositionwiseFeedForward, EncoderLayer, SequentialEncoder, Encoder, MultiHeadCrossAttention, DecoderLayer,
SequentialDecoder, Decoder, Transformer), implemented on the b200ml kernels.  ``from transformer import Transformer``
works as in pytorch_machine_translator.py:9."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from b200ml.models.transformer import (  # noqa: E402,F401
    Decoder, DecoderLayer, Encoder, EncoderLayer, LayerNormalization, MultiHeadAttention, MultiHeadCrossAttention,
    PositionalEncoding, PositionwiseFeedForward, SentenceEmbedding, SequentialDecoder, SequentialEncoder, Transformer,
    get_device, scaled_dot_product)
