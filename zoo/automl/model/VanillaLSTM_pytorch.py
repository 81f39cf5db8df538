"""VanillaLSTM (reference: pyzoo/zoo/automl/model/VanillaLSTM_pytorch.py:29-270; keras twin VanillaLSTM.py).
`LSTMModel` keeps the reference's sub-module names (`lstm`, `fc`) and init, so its `state_dict` files are
interchangeable; the sequence loop runs on kernel K6."""
from __future__ import annotations

import torch.nn as nn

from ...ops.rnn import run as rnn_run
from .torch_base import TorchForecastModel


class LSTMModel(nn.Module):
    def __init__(self, input_dim, hidden_dim, layer_num, dropout, output_dim):
        super().__init__()
        self.hidden_dim, self.layer_num = hidden_dim, layer_num
        self.lstm = nn.LSTM(input_dim, hidden_dim, layer_num, dropout=dropout if layer_num > 1 else 0.0, batch_first=True)
        self.fc = nn.Linear(hidden_dim, output_dim)
        for name, param in self.lstm.named_parameters():
            if "bias" in name:
                nn.init.constant_(param, 0.0)
            elif "weight_ih" in name:
                nn.init.xavier_normal_(param)
            elif "weight_hh" in name:
                nn.init.orthogonal_(param)

    def forward(self, input_seq):
        lstm_out, _ = rnn_run(self.lstm, input_seq.float())
        return self.fc(lstm_out[:, -1, :])


class TwoLayerLSTM(nn.Module):
    """The keras VanillaLSTM: LSTM(lstm_1_units) -> Dropout -> LSTM(lstm_2_units) -> Dropout -> Dense."""

    def __init__(self, input_dim, units1, drop1, units2, drop2, output_dim):
        super().__init__()
        self.lstm1, self.lstm2 = nn.LSTM(input_dim, units1, batch_first=True), nn.LSTM(units1, units2, batch_first=True)
        self.drop1, self.drop2 = nn.Dropout(drop1), nn.Dropout(drop2)
        self.fc = nn.Linear(units2, output_dim)

    def forward(self, x):
        h, _ = rnn_run(self.lstm1, x.float())
        h, _ = rnn_run(self.lstm2, self.drop1(h))
        return self.fc(self.drop2(h[:, -1]))


class VanillaLSTMPytorch(TorchForecastModel):
    def _get_configs(self, config):
        super()._get_configs(config)
        self.hidden_dim = int(config.get("hidden_dim", 32))
        self.layer_num = int(config.get("layer_num", 2))
        self.dropout = float(config.get("dropout", 0.2))

    def _build(self, config):
        if "lstm_1_units" in config:     # the keras recipe keys
            return TwoLayerLSTM(self.feature_num, int(config["lstm_1_units"]), float(config.get("dropout_1", 0.2)),
                                int(config.get("lstm_2_units", 16)), float(config.get("dropout_2", 0.2)), self.future_seq_len)
        return LSTMModel(self.feature_num, self.hidden_dim, self.layer_num, self.dropout, self.future_seq_len)

    def _get_required_parameters(self):
        return set()

    def _get_optional_parameters(self):
        return {"hidden_dim", "layer_num", "dropout", "lr", "batch_size", "lstm_1_units", "dropout_1", "lstm_2_units", "dropout_2"}


VanillaLSTM = VanillaLSTMPytorch
