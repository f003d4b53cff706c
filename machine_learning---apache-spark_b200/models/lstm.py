"""LSTM text classifier of the reference (distributed_lstm.py:110-135, pytorch_lstm.py:94-119):
embedding → 2-layer LSTM (batch_first, inter-layer dropout 0.5) → Linear on every time step.

Same state_dict keys (embedding.weight, lstm.{weight,bias}_{ih,hh}_l{0,1}, fc_out.{weight,bias}; gate order i,f,g,o).
Constructor quirks fixed per SURVEY §9: the arguments are honoured (Q6: the distributed copy ignores them),
``padding_idx`` is explicit (default 0 = '<pad>'), and fc_out takes ``hidden_size`` inputs instead of a literal 32 (Q8).
The recurrence runs in one persistent kernel per layer stack (ops/lstm.py) instead of T×layers cell launches.
"""
from __future__ import annotations

import torch
from torch import nn

from ..ops import lstm as _lstm
from ..ops.embedding import embedding as _embedding
from ..ops.gemm import linear as _linear
from ..utils.rng import new_site


class LSTM(nn.Module):
    def __init__(self, vocab_size, embedding_dim, hidden_size, output_size, num_layers=2, padding_idx=0,
                 dropout: float = 0.5):
        super().__init__()
        self.embedding = nn.Embedding(num_embeddings=vocab_size, embedding_dim=embedding_dim, padding_idx=padding_idx)
        self.embedding.weight._b200_sparse_rows = True   # row-sparse gradient: the DP engine skips untouched rows
        self.lstm = nn.LSTM(input_size=embedding_dim, hidden_size=hidden_size, num_layers=num_layers, batch_first=True,
                            dropout=dropout)
        self.fc_out = nn.Linear(hidden_size, output_size)
        self.dropout = nn.Dropout(0.5)  # defined (and unused) in pytorch_lstm.py:109; no parameters
        self.hidden_size, self.num_layers = hidden_size, num_layers
        self._site = new_site()

    def forward(self, input_seq, hidden_in, mem_in):
        emb = _embedding(input_seq, self.embedding.weight, padding_idx=self.embedding.padding_idx,
                         out_dtype=torch.float32)
        weights = [getattr(self.lstm, n) for n in self.lstm._flat_weights_names]
        output, hidden_out, mem_out = _lstm.lstm_forward(emb, hidden_in, mem_in, weights, self.num_layers,
                                                         self.lstm.dropout if self.training else 0.0, self._site)
        return _linear(output, self.fc_out.weight, self.fc_out.bias).float(), hidden_out, mem_out
