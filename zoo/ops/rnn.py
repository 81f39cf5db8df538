"""Recurrent layers (kernel K6 face).

LSTM / GRU / SimpleRNN modules with the call shape the zoo keras layers need (`return_sequences`,
`go_backwards`, `inner_activation`), batch first.  Weights are laid out like ``torch.nn.LSTM/GRU``
(gate order i,f,g,o / r,z,n) so state_dicts of the reference's PyTorch AutoML models
(`VanillaLSTM_pytorch.py:29-56`, `Seq2Seq_pytorch.py:31-101`) load unchanged.  On GPU the sequence
loop runs in the persistent kernels of ``csrc/ops/rnn.cu``; on CPU in torch (cuDNN-free cell loop when
the zoo keras default `hard_sigmoid` gate activation is requested, which torch's fused RNN lacks).
"""
from __future__ import annotations

import torch
import torch.nn as nn


def hard_sigmoid(x):
    return torch.clamp(0.2 * x + 0.5, 0.0, 1.0)


def cell_loop(rnn: nn.Module, x, state=None, hard=False):
    """Reference implementation of nn.LSTM / nn.GRU (unidirectional, batch-first) as an explicit loop
    with a selectable gate activation: the CPU path and the oracle of the kernel tests."""
    gate = hard_sigmoid if hard else torch.sigmoid
    is_lstm = isinstance(rnn, nn.LSTM)
    B, T, _ = x.shape
    H = rnn.hidden_size
    out, hs, cs = x, [], []
    for l in range(rnn.num_layers):
        w_ih, w_hh = getattr(rnn, f"weight_ih_l{l}"), getattr(rnn, f"weight_hh_l{l}")
        b_ih = getattr(rnn, f"bias_ih_l{l}") if rnn.bias else None
        b_hh = getattr(rnn, f"bias_hh_l{l}") if rnn.bias else None
        h = state[0][l] if (is_lstm and state is not None) else (state[l] if state is not None else x.new_zeros(B, H))
        c = state[1][l] if (is_lstm and state is not None) else x.new_zeros(B, H)
        gx = torch.nn.functional.linear(out, w_ih, b_ih)
        ys = []
        for t in range(T):
            gh = torch.nn.functional.linear(h, w_hh, b_hh)
            if is_lstm:
                g = gx[:, t] + gh
                i, f, gg, o = g.chunk(4, -1)
                c = gate(f) * c + gate(i) * torch.tanh(gg)
                h = gate(o) * torch.tanh(c)
            else:
                xr, xz, xn = gx[:, t].chunk(3, -1)
                hr, hz, hn = gh.chunk(3, -1)
                r, z = gate(xr + hr), gate(xz + hz)
                n = torch.tanh(xn + r * hn)
                h = (1 - z) * n + z * h
            ys.append(h)
        out = torch.stack(ys, 1)
        hs.append(h)
        cs.append(c)
        if rnn.dropout > 0 and l + 1 < rnn.num_layers and rnn.training:
            out = torch.nn.functional.dropout(out, rnn.dropout, True)
    hN = torch.stack(hs)
    return (out, (hN, torch.stack(cs))) if is_lstm else (out, hN)


class _Recurrent(nn.Module):
    cell = None

    def __init__(self, input_size, hidden_size, num_layers=1, dropout=0.0, go_backwards=False, inner_activation="sigmoid"):
        super().__init__()
        self.input_size, self.hidden_size, self.num_layers = int(input_size), int(hidden_size), int(num_layers)
        self.go_backwards = go_backwards
        if inner_activation not in ("sigmoid", "hard_sigmoid"):
            raise ValueError("inner_activation must be 'sigmoid' or 'hard_sigmoid'")
        self.hard = inner_activation == "hard_sigmoid"
        self.rnn = self.cell(self.input_size, self.hidden_size, num_layers=self.num_layers, batch_first=True,
                             dropout=dropout if num_layers > 1 else 0.0)

    def forward(self, x, state=None, return_sequences=False, return_state=False):
        x = x.float()
        if self.go_backwards:
            x = x.flip(1)
        out, st = run(self.rnn, x, state, self.hard)
        y = out if return_sequences else out[:, -1]
        return (y, st) if return_state else y


def run(rnn: nn.Module, x, state=None, hard=False):
    """(outputs [B,T,H], final_state).  CUDA + covered module -> persistent kernels (a missing native
    library is an error there, never a silent fallback); otherwise torch."""
    if x.is_cuda:
        from .lstm_persistent import maybe_run

        r = maybe_run(rnn, x, state, hard)
        if r is not None:
            return r
    if hard:
        return cell_loop(rnn, x, state, True)
    return rnn(x, state)


class LSTM(_Recurrent):
    cell = nn.LSTM


class GRU(_Recurrent):
    cell = nn.GRU


class SimpleRNN(nn.Module):
    """Elman RNN (zoo keras SimpleRNN): h_t = act(W x_t + U h_{t-1} + b).  Short sequences of small width:
    runs through torch's fused RNN."""

    def __init__(self, input_size, hidden_size, activation="tanh", go_backwards=False):
        super().__init__()
        if activation not in ("tanh", "relu"):
            raise ValueError("SimpleRNN supports tanh and relu")
        self.rnn = nn.RNN(input_size, hidden_size, nonlinearity=activation, batch_first=True)
        self.go_backwards = go_backwards
        self.hidden_size = hidden_size

    def forward(self, x, state=None, return_sequences=False, return_state=False):
        x = x.float()
        if self.go_backwards:
            x = x.flip(1)
        out, st = self.rnn(x, state)
        y = out if return_sequences else out[:, -1]
        return (y, st) if return_state else y
