"""MTNet: memory time-series network (reference: pyzoo/zoo/automl/model/MTNet_keras.py:54-606; Chang et al. 2018).

Input `[batch, (long_num + 1) * time_step, features]`: the first `long_num` blocks are the long-term memory,
the last block is the short-term query.  Every block goes through an encoder -- a causal convolution over
time of height `cnn_height` spanning all features (kernel K7) followed by a GRU (kernel K6) with attention
over its states; memory blocks are encoded twice (input memory `m_i`, output memory `c_i`), the query once
(`u`); attention `p = softmax(m_i . u)` mixes the output memory, `[sum p_i c_i ; u]` feeds a linear head, and
an autoregressive linear term over the last `ar_size` target values is added (MTNet_keras.py:382-432).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ...ops.rnn import run as rnn_run
from ...ops.tcn import causal_conv1d
from .torch_base import TorchForecastModel


class _Encoder(nn.Module):
    def __init__(self, feature_num, cnn_height, cnn_hid_size, rnn_hid_sizes, cnn_dropout, rnn_dropout):
        super().__init__()
        self.conv = nn.Conv1d(feature_num, cnn_hid_size, cnn_height)
        self.cnn_dropout, self.rnn_dropout = cnn_dropout, rnn_dropout
        sizes = [cnn_hid_size] + list(rnn_hid_sizes)
        self.grus = nn.ModuleList([nn.GRU(a, b, batch_first=True) for a, b in zip(sizes[:-1], sizes[1:])])
        self.attn = nn.Linear(sizes[-1], 1)

    def forward(self, x):                                   # x: [N, T, F]
        h = causal_conv1d(x.transpose(1, 2).contiguous(), self.conv.weight, self.conv.bias, 1, True, None, self.cnn_dropout, self.training)
        h = h.transpose(1, 2)
        for g in self.grus:
            h, _ = rnn_run(g, h)
            h = torch.nn.functional.dropout(h, self.rnn_dropout, self.training)
        w = torch.softmax(self.attn(h), dim=1)              # attention over time
        return (w * h).sum(1)


class MTNetModule(nn.Module):
    def __init__(self, feature_num, output_dim, time_step, long_num, ar_size, cnn_height, cnn_hid_size, rnn_hid_sizes, cnn_dropout, rnn_dropout):
        super().__init__()
        self.time_step, self.long_num, self.ar_size = time_step, long_num, ar_size
        mk = lambda: _Encoder(feature_num, cnn_height, cnn_hid_size, rnn_hid_sizes, cnn_dropout, rnn_dropout)
        self.enc_m, self.enc_c, self.enc_u = mk(), mk(), mk()
        self.head = nn.Linear(2 * rnn_hid_sizes[-1], output_dim)
        self.ar = nn.Linear(ar_size, output_dim)

    def forward(self, x):
        B, L, F = x.shape
        T, n = self.time_step, self.long_num
        if L != (n + 1) * T:
            raise ValueError(f"MTNet expects past_seq_len == (long_num + 1) * time_step = {(n + 1) * T}, got {L}")
        x = x.float()
        mem = x[:, :n * T].reshape(B * n, T, F)
        q = x[:, n * T:]
        m = self.enc_m(mem).reshape(B, n, -1)
        c = self.enc_c(mem).reshape(B, n, -1)
        u = self.enc_u(q)
        p = torch.softmax(torch.bmm(m, u.unsqueeze(2)), dim=1)          # [B, n, 1]
        o = (p * c).sum(1)
        return self.head(torch.cat([o, u], 1)) + self.ar(q[:, -self.ar_size:, 0])


class MTNet(TorchForecastModel):
    def _get_configs(self, config):
        super()._get_configs(config)
        self.time_step = int(config.get("time_step", 3))
        self.long_num = int(config.get("long_num", 3))
        self.ar_size = int(config.get("ar_size", config.get("ar_window", 2)))
        self.cnn_height = int(config.get("cnn_height", config.get("filter_size", 2)))
        self.cnn_hid_size = int(config.get("cnn_hid_size", 32))
        r = config.get("rnn_hid_sizes", [16, 32])
        self.rnn_hid_sizes = [int(v) for v in (r if isinstance(r, (list, tuple)) else [r])]
        d = float(config.get("dropout", 0.2))
        self.cnn_dropout, self.rnn_dropout = float(config.get("cnn_dropout", d)), float(config.get("rnn_dropout", d))

    def _build(self, config):
        if self.ar_size > self.time_step:
            raise ValueError(f"ar_size {self.ar_size} must not exceed time_step {self.time_step}")
        return MTNetModule(self.feature_num, self.future_seq_len, self.time_step, self.long_num, self.ar_size, self.cnn_height,
                           self.cnn_hid_size, self.rnn_hid_sizes, self.cnn_dropout, self.rnn_dropout)

    def _get_required_parameters(self):
        return {"time_step", "long_num"}

    def _get_optional_parameters(self):
        return {"ar_size", "cnn_height", "cnn_hid_size", "rnn_hid_sizes", "cnn_dropout", "rnn_dropout", "lr", "batch_size"}


MTNetKeras = MTNet
