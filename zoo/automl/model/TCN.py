"""TCN forecaster (north-star extension; no TCN in the reference snapshot): a stack of
weight-normalised dilated causal conv blocks (kernel K7) and a linear head on the last time step."""
from __future__ import annotations

import torch.nn as nn

from ...ops.tcn import TemporalConvNet
from .torch_base import TorchForecastModel


class TCNModule(nn.Module):
    def __init__(self, feature_num, output_dim, num_channels, kernel_size, dropout):
        super().__init__()
        self.tcn = TemporalConvNet(feature_num, num_channels, kernel_size, dropout)
        self.linear = nn.Linear(num_channels[-1], output_dim)
        nn.init.normal_(self.linear.weight, 0.0, 0.01)

    def forward(self, x):                        # [B, T, F]
        y = self.tcn(x.float().transpose(1, 2).contiguous())
        return self.linear(y[:, :, -1])


class TCNPytorch(TorchForecastModel):
    def _get_configs(self, config):
        super()._get_configs(config)
        self.hidden_units, self.levels = int(config.get("hidden_units", 30)), int(config.get("levels", 4))
        ch = config.get("num_channels")
        self.num_channels = [int(c) for c in ch] if ch else [self.hidden_units] * self.levels
        self.kernel_size, self.dropout = int(config.get("kernel_size", 3)), float(config.get("dropout", 0.1))

    def _build(self, config):
        return TCNModule(self.feature_num, self.future_seq_len, self.num_channels, self.kernel_size, self.dropout)


TCN = TCNPytorch
