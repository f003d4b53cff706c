"""Multi-layer LSTM recurrence — K9 (csrc/kernels/lstm.cu).  Reference: nn.LSTM(E, H, num_layers=2, batch_first=True,
dropout=0.5) inside the text classifier (distributed_lstm.py:118-119,131); gate order i, f, g, o."""
from __future__ import annotations

import torch

from .. import _native


def lstm_reference(x, h0, c0, weights, num_layers: int, dropout_p: float):
    """Plain PyTorch restatement of the stacked LSTM (differentiable).  x: [B,T,E]; h0/c0: [L,B,H];
    weights: [w_ih_l0, w_hh_l0, b_ih_l0, b_hh_l0, w_ih_l1, ...]."""
    B, T, _ = x.shape
    inp = x.float()
    hs, cs = [], []
    for layer in range(num_layers):
        w_ih, w_hh, b_ih, b_hh = weights[4 * layer:4 * layer + 4]
        h, c = h0[layer].float(), c0[layer].float()
        pre = inp @ w_ih.float().t() + b_ih.float() + b_hh.float()   # input projection hoisted out of the recurrence
        outs = []
        for t in range(T):
            g = pre[:, t] + h @ w_hh.float().t()
            i, f, gg, o = g.chunk(4, dim=-1)
            c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
            h = torch.sigmoid(o) * torch.tanh(c)
            outs.append(h)
        inp = torch.stack(outs, dim=1)
        if dropout_p > 0 and layer < num_layers - 1:
            inp = torch.nn.functional.dropout(inp, dropout_p, training=True)
        hs.append(h)
        cs.append(c)
    return inp, torch.stack(hs), torch.stack(cs)


def lstm_forward(x, h0, c0, weights, num_layers: int, dropout_p: float = 0.0, site: int = 0):
    """Returns (output [B,T,H], h_n [L,B,H], c_n [L,B,H])."""
    if _native.use_native(x):
        from . import lstm_native

        return lstm_native.lstm_forward(x, h0, c0, weights, num_layers, dropout_p, site)
    return lstm_reference(x, h0, c0, weights, num_layers, dropout_p)
