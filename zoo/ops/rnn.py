"""Recurrent layers (kernel K6 face).

LSTM / GRU modules with the call shape the zoo keras layers need (`return_sequences`), batch
first.  Weights are laid out like ``torch.nn.LSTM/GRU`` (gate order i,f,g,o / r,z,n) so
state_dicts of the reference's PyTorch AutoML models (`VanillaLSTM_pytorch.py:29-56`,
`Seq2Seq_pytorch.py:31-101`) load unchanged.  On GPU the sequence loop runs in the persistent
recurrent kernels of ``csrc/ops/rnn.cu`` when they are available for the shape; otherwise
cuDNN through torch.
"""
from __future__ import annotations

import torch
import torch.nn as nn


class _Recurrent(nn.Module):
    cell = None

    def __init__(self, input_size, hidden_size, num_layers=1, dropout=0.0, go_backwards=False):
        super().__init__()
        self.input_size, self.hidden_size, self.num_layers = int(input_size), int(hidden_size), int(num_layers)
        self.go_backwards = go_backwards
        self.rnn = self.cell(self.input_size, self.hidden_size, num_layers=self.num_layers, batch_first=True,
                             dropout=dropout if num_layers > 1 else 0.0)

    def forward(self, x, state=None, return_sequences=False, return_state=False):
        x = x.float()
        if self.go_backwards:
            x = x.flip(1)
        from . import rnn_kernels

        out, st = rnn_kernels.run(self.rnn, x, state)
        y = out if return_sequences else out[:, -1]
        return (y, st) if return_state else y


class LSTM(_Recurrent):
    cell = nn.LSTM


class GRU(_Recurrent):
    cell = nn.GRU
