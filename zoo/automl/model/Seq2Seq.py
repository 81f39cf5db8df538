"""`LSTMSeq2Seq`, the Keras-recipe encoder / decoder forecaster (reference: pyzoo/zoo/automl/model/Seq2Seq.py:30-330).

What the reference's model does and this one keeps: ONE LSTM layer of `latent_dim` (128) units on each side with `dropout` (0.2) on
the layer inputs; the encoder's final (h, c) start the decoder; training is teacher-forced on the LAGGED target (first decoder
input = the last observed target value, Seq2Seq.py:150-163); prediction decodes step by step feeding its own output back
(Seq2Seq.py:120-147); several target columns (`y` of shape [n, future_seq_len, target_col_num]); RMSprop(lr), batch_size 64;
Monte-Carlo dropout (`mc=True`) keeps the dropouts active while predicting.  The recurrences run on the LSTM kernel (K6).
"""
import numpy as np
import torch
import torch.nn as nn

from ...ops.rnn import run as rnn_run
from .torch_base import TorchForecastModel


class _KerasSeq2Seq(nn.Module):
    def __init__(self, feature_num, target_col_num, latent_dim, dropout, future_seq_len):
        super().__init__()
        self.encoder_lstm = nn.LSTM(feature_num, latent_dim, batch_first=True)
        self.decoder_lstm = nn.LSTM(target_col_num, latent_dim, batch_first=True)
        self.decoder_dense = nn.Linear(latent_dim, target_col_num)
        self.enc_drop, self.dec_drop = nn.Dropout(dropout), nn.Dropout(dropout)
        self.target_col_num, self.future_seq_len = target_col_num, future_seq_len

    def encode(self, x):
        _, state = rnn_run(self.encoder_lstm, self.enc_drop(x.float()))
        return state

    def forward(self, x, decoder_inputs=None):
        """Teacher-forced when `decoder_inputs` [B, future_seq_len, target_col_num] is given, else step-wise decoding."""
        state = self.encode(x)
        if decoder_inputs is not None:
            out, _ = rnn_run(self.decoder_lstm, self.dec_drop(decoder_inputs.float()), state)
            return self.decoder_dense(out)
        step = x[:, -1:, :self.target_col_num].float()
        outs = []
        for _ in range(self.future_seq_len):
            out, state = rnn_run(self.decoder_lstm, self.dec_drop(step), state)
            step = self.decoder_dense(out)
            outs.append(step)
        return torch.cat(outs, dim=1)


class LSTMSeq2Seq(TorchForecastModel):
    def _get_configs(self, config):
        super()._get_configs(config)
        self.batch_size = int(config.get("batch_size", 64))
        self.latent_dim = int(config.get("latent_dim", 128))
        self.dropout = float(config.get("dropout", 0.2))
        self.target_col_num = int(config.get("target_col_num", getattr(self, "target_col_num", 1)))

    def _prepare_y(self, y):
        y = np.asarray(y, dtype=np.float32)
        while y.ndim < 3:                       # [n] -> [n, 1, 1];  [n, future] -> [n, future, 1]  (Seq2Seq.py:171-179)
            y = y[..., None] if y.ndim == 2 else y[:, None]
        return y

    def _ensure_model(self, x, config):
        if self.model is None and "_y_shape" in self.__dict__:
            self.future_seq_len, self.target_col_num = int(self._y_shape[1]), int(self._y_shape[2])
        super()._ensure_model(x, config)

    def fit_eval(self, x, y, validation_data=None, mc=False, verbose=0, **config):
        self._y_shape = self._prepare_y(y).shape
        return super().fit_eval(x, y, validation_data, mc, verbose, **config)

    def _build(self, config):
        return _KerasSeq2Seq(self.feature_num, self.target_col_num, self.latent_dim, self.dropout, self.future_seq_len)

    def _make_optimizer(self):
        return torch.optim.RMSprop(self.model.parameters(), lr=self.lr, alpha=0.9, eps=1e-7)

    def _train_loss(self, xb, yb):
        first = xb[:, -1:, :self.target_col_num].float()
        decoder_inputs = torch.cat([first, yb[:, :-1]], dim=1)       # the target, one step behind
        return self.criterion(self.model(xb, decoder_inputs), yb)

    def predict(self, x, mc=False):
        if self.model is None:
            raise Exception("Needs to call fit_eval or restore first before calling predict")
        out = self._forward_all(x, train_mode=mc)                    # [n, future_seq_len, target_col_num]
        return out[..., 0] if self.target_col_num == 1 else out

    def evaluate(self, x, y, metric=None):
        from ..common.metrics import Evaluator

        yp, yt = self.predict(x), self._prepare_y(y)
        return [Evaluator.evaluate(m, yt.reshape(len(yt), -1), np.asarray(yp).reshape(len(yt), -1), multioutput="uniform_average")
                for m in (metric or ["mse"])]

    def save(self, model_path, config_path):
        self.config = dict(self.config, target_col_num=self.target_col_num, latent_dim=self.latent_dim, dropout=self.dropout)
        super().save(model_path, config_path)

    def _get_required_parameters(self):
        return set()

    def _get_optional_parameters(self):
        return {"past_seq_len", "latent_dim", "dropout", "metric", "lr", "epochs", "batch_size"}
