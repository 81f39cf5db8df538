"""Standalone forecasters (reference: pyzoo/zoo/zouwu/model/forecast.py:26-166): `LSTMForecaster`,
`MTNetForecaster` -- TFPark `KerasModel` subclasses there, i.e. `fit / evaluate / predict` on ndarrays with
compile-time loss / optimizer / metrics.  Here they are zoo `KerasNet`s over the torch models (kernels K6 / K7),
so the same three calls run through the training engine.  `TCNForecaster` is the north-star extension."""
from __future__ import annotations

import numpy as np

from ...automl.model.MTNet import MTNetModule
from ...automl.model.TCN import TCNModule
from ...automl.model.VanillaLSTM_pytorch import TwoLayerLSTM
from ...pipeline.api.keras.engine.topology import KerasNet


class Forecaster(KerasNet):
    """Base: builds `self.net`, compiles with (loss, optimizer, metrics)."""

    def __init__(self, net, loss="mse", optimizer="adam", metrics=None, lr=0.001, uncertainty=False):
        super().__init__()
        self.net = net
        self.uncertainty = uncertainty
        from ...pipeline.api.keras.optimizers import Adam

        opt = Adam(lr=lr) if (isinstance(optimizer, str) and optimizer.lower() == "adam") else optimizer
        self.compile(optimizer=opt, loss=loss, metrics=list(metrics or ["mse"]))

    def forward(self, x):
        return self.net(x)

    def fit(self, x, y=None, validation_data=None, batch_size=32, epochs=1, nb_epoch=None, distributed=False):
        y = None if y is None else np.asarray(y, dtype=np.float32).reshape(len(y), -1)
        if validation_data is not None:
            validation_data = (validation_data[0], np.asarray(validation_data[1], dtype=np.float32).reshape(len(validation_data[1]), -1))
        return super().fit(np.asarray(x, dtype=np.float32), y, batch_size=batch_size, nb_epoch=nb_epoch or epochs,
                           validation_data=validation_data)

    def evaluate(self, x, y=None, batch_size=32, batch_per_thread=None, distributed=False):
        y = None if y is None else np.asarray(y, dtype=np.float32).reshape(len(y), -1)
        return super().evaluate(np.asarray(x, dtype=np.float32), y, batch_size=batch_size)

    def predict(self, x, batch_per_thread=None, distributed=False, batch_size=None):
        return super().predict(np.asarray(x, dtype=np.float32), batch_size=batch_size or 1024)

    def predict_with_uncertainty(self, x, n_iter=100):
        import torch

        dev = next(self.parameters()).device
        xt = torch.as_tensor(np.asarray(x, dtype=np.float32)).to(dev)
        self.train()                      # MC dropout: keep dropout active
        with torch.no_grad():
            runs = torch.stack([self.net(xt).float() for _ in range(n_iter)])
        self.eval()
        return runs.mean(0).cpu().numpy(), runs.std(0).cpu().numpy()


class LSTMForecaster(Forecaster):
    """Vanilla LSTM forecaster: x `[n, lookback, feature_dim]` -> `[n, target_dim]` (forecast.py:49-105)."""

    def __init__(self, target_dim=1, feature_dim=1, lstm_1_units=16, dropout_1=0.2, lstm_2_units=8, dropout_2=0.2, metric="mean_squared_error",
                 lr=0.001, loss="mse", uncertainty: bool = False, optimizer="adam"):
        self.target_dim, self.feature_dim = target_dim, feature_dim
        self.model_config = dict(lr=lr, lstm_1_units=lstm_1_units, dropout_1=dropout_1, lstm_2_units=lstm_2_units, dropout_2=dropout_2,
                                 metric=metric, feature_num=feature_dim, output_dim=target_dim)
        net = TwoLayerLSTM(feature_dim, lstm_1_units, dropout_1, lstm_2_units, dropout_2, target_dim)
        super().__init__(net, loss=loss, optimizer=optimizer, metrics=["mse"], lr=lr, uncertainty=uncertainty)


class MTNetForecaster(Forecaster):
    """MTNet forecaster: x `[n, (long_series_num + 1) * series_length, feature_dim]` (forecast.py:108-166);
    `preprocess_input` splits it into the reference's (long-term, short-term) pair."""

    def __init__(self, target_dim=1, feature_dim=1, long_series_num=1, series_length=1, ar_window_size=1, cnn_height=1, cnn_hid_size=32,
                 rnn_hid_sizes=(16, 32), lr=0.001, loss="mae", cnn_dropout=0.2, rnn_dropout=0.2, metric="mean_absolute_error",
                 uncertainty: bool = False, optimizer="adam"):
        self.long_series_num, self.series_length = long_series_num, series_length
        self.model_config = dict(feature_num=feature_dim, output_dim=target_dim, long_num=long_series_num, time_step=series_length,
                                 ar_window=ar_window_size, cnn_height=cnn_height, cnn_hid_size=cnn_hid_size, rnn_hid_sizes=list(rnn_hid_sizes),
                                 cnn_dropout=cnn_dropout, rnn_dropout=rnn_dropout, lr=lr, metric=metric)
        net = MTNetModule(feature_dim, target_dim, series_length, long_series_num, ar_window_size, cnn_height, cnn_hid_size,
                          list(rnn_hid_sizes), cnn_dropout, rnn_dropout)
        super().__init__(net, loss=loss, optimizer=optimizer, metrics=["mae"], lr=lr, uncertainty=uncertainty)

    def preprocess_input(self, x):
        """-> [long_term `[n, long_series_num, series_length, feature]`, short_term `[n, series_length, feature]`]."""
        x = np.asarray(x)
        n, T = self.long_series_num, self.series_length
        if x.shape[1] != (n + 1) * T:
            raise ValueError(f"x.shape[1] should be (long_series_num + 1) * series_length = {(n + 1) * T}, got {x.shape[1]}")
        return [x[:, :n * T].reshape(len(x), n, T, x.shape[2]), x[:, n * T:]]


class TCNForecaster(Forecaster):
    def __init__(self, target_dim=1, feature_dim=1, num_channels=(30, 30, 30, 30), kernel_size=3, dropout=0.1, lr=0.001, loss="mse",
                 uncertainty: bool = False, optimizer="adam"):
        net = TCNModule(feature_dim, target_dim, list(num_channels), kernel_size, dropout)
        super().__init__(net, loss=loss, optimizer=optimizer, metrics=["mse"], lr=lr, uncertainty=uncertainty)
