"""Shared training loop of the torch forecasting models (VanillaLSTM / Seq2Seq / MTNet / TCN).

Reference shape: `VanillaLSTMPytorch` (pyzoo/zoo/automl/model/VanillaLSTM_pytorch.py:59-270) -- `fit_eval`
builds the model on first call, trains `epochs` epochs with Adam + MSE, returns the validation metric;
`predict(mc=True)` keeps dropout on; `predict_with_uncertainty` = MC-dropout mean/std over `n_iter` passes;
`save` writes the `state_dict` (+ config JSON), `restore` rebuilds from the config and loads it.
The recurrent / convolutional cores run on kernels K6 / K7 when the model sits on a GPU.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from ..common.metrics import Evaluator
from ..common.util import save_config
from .abstract import BaseModel


def default_device():
    return torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")


class TorchForecastModel(BaseModel):
    saved_config_keys = ()

    def __init__(self, check_optional_config=True, future_seq_len=1, device=None):
        self.check_optional_config, self.future_seq_len = check_optional_config, int(future_seq_len)
        self.model, self.optimizer, self.criterion = None, None, nn.MSELoss()
        self.feature_num, self.metric, self.batch_size, self.lr = None, "mse", 1024, 1e-3
        self.device = torch.device(device) if device is not None else default_device()
        self.config = {}

    # subclasses: read hyper-parameters and build the nn.Module
    def _get_configs(self, config):
        self.metric = config.get("metric", "mse")
        self.batch_size = int(config.get("batch_size", 1024))
        self.lr = float(config.get("lr", 1e-3))

    def _build(self, config):  # pragma: no cover - interface
        raise NotImplementedError

    def _prepare_y(self, y):
        y = np.asarray(y, dtype=np.float32)
        return y.reshape(len(y), -1)

    def _ensure_model(self, x, config):
        if self.model is None:
            self.feature_num = int(x.shape[2])
            self.past_seq_len = int(x.shape[1])
            self._get_configs(config)
            self.config = {k: config[k] for k in config if isinstance(config[k], (int, float, str, bool, list, np.integer, np.floating))}
            self.model = self._build(config).to(self.device)
            self.optimizer = self._make_optimizer()

    def _make_optimizer(self):
        return torch.optim.Adam(self.model.parameters(), lr=self.lr)

    def _train_loss(self, xb, yb):
        return self.criterion(self.model(xb).reshape(yb.shape), yb)

    def _batches(self, x, y=None, shuffle=False):
        n = len(x)
        order = np.random.permutation(n) if shuffle else np.arange(n)
        xt = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        yt = None if y is None else torch.from_numpy(y)
        for s in range(0, n, self.batch_size):
            idx = torch.from_numpy(order[s:s + self.batch_size])
            yield xt[idx].to(self.device, non_blocking=True), (None if yt is None else yt[idx].to(self.device, non_blocking=True))

    def fit_eval(self, x, y, validation_data=None, mc=False, verbose=0, **config):
        self._ensure_model(x, config)
        y2 = self._prepare_y(y)
        epochs = int(config.get("epochs", 1))
        for _ in range(epochs):
            self.model.train()
            for xb, yb in self._batches(x, y2, shuffle=True):
                self.optimizer.zero_grad(set_to_none=True)
                loss = self._train_loss(xb, yb)
                loss.backward()
                self.optimizer.step()
        vx, vy = validation_data if validation_data is not None else (x, y)
        return float(self.evaluate(vx, vy, [self.metric])[0])

    def _forward_all(self, x, train_mode=False):
        self.model.train(train_mode)
        outs = []
        with torch.no_grad():
            for xb, _ in self._batches(x):
                outs.append(self.model(xb).float().cpu())
        self.model.eval()
        return torch.cat(outs).numpy()

    def predict(self, x, mc=False):
        if self.model is None:
            raise Exception("Needs to call fit_eval or restore first before calling predict")
        out = self._forward_all(x, train_mode=mc)
        return out.reshape(len(x), -1) if self.future_seq_len > 1 else out.reshape(len(x), -1)[:, :1]

    def predict_with_uncertainty(self, x, n_iter=100):
        runs = np.stack([self.predict(x, mc=True) for _ in range(n_iter)])
        return runs.mean(0), runs.std(0)

    def evaluate(self, x, y, metric=None):
        metric = metric or ["mse"]
        yp = self.predict(x)
        yt = self._prepare_y(y)
        return [Evaluator.evaluate(m, yt, yp.reshape(yt.shape), multioutput="uniform_average") for m in metric]

    def save(self, model_path, config_path):
        torch.save(self.model.state_dict(), model_path)
        cfg = dict(self.config)
        cfg.update({"future_seq_len": self.future_seq_len, "feature_num": self.feature_num, "past_seq_len": self.past_seq_len,
                    "model_class": type(self).__name__, "metric": self.metric})
        save_config(config_path, cfg)

    def restore(self, model_path, **config):
        self.future_seq_len = int(config.get("future_seq_len", self.future_seq_len))
        self.feature_num, self.past_seq_len = int(config["feature_num"]), int(config.get("past_seq_len", 2))
        self._get_configs(config)
        self.config = dict(config)
        self.model = self._build(config).to(self.device)
        self.model.load_state_dict(torch.load(model_path, map_location=self.device))
        self.optimizer = self._make_optimizer()
        return self
