"""AnomalyDetector (reference: pyzoo/zoo/models/anomalydetection/anomaly_detector.py:30-183,
zoo/.../models/anomalydetection/AnomalyDetector.scala:40-173): stacked LSTM (+Dropout) -> Dense(1) regressor over
unrolled windows; anomalies = the `anomaly_size` largest |truth - prediction|.  LSTMs run on kernel K6."""
from __future__ import annotations

import numpy as np

from ...pipeline.api.keras.engine.topology import Sequential
from ...pipeline.api.keras.layers import LSTM, Dense, Dropout, InputLayer
from ..common.zoo_model import KerasZooModel


class FeatureLableIndex:
    """One unrolled sample: feature window [unroll_length, n_features], label, index of the label row."""

    def __init__(self, feature, label, index, bigdl_type="float"):
        self.feature, self.label, self.index = feature, label, index

    def __reduce__(self):
        return FeatureLableIndex, (self.feature, self.label, self.index)

    def __str__(self):
        return f"value: {self.feature}, label:{self.label}, index: {self.index}"


class AnomalyDetector(KerasZooModel):
    def __init__(self, feature_shape, hidden_layers=(8, 32, 15), dropouts=(0.2, 0.2, 0.2), **kwargs):
        kwargs.pop("bigdl_type", None)
        if kwargs:
            raise TypeError("Wrong arguments for AnomalyDetector: " + str(kwargs))
        assert len(hidden_layers) == len(dropouts), "sizes of dropouts and hidden_layers should be equal"
        super().__init__()
        self.feature_shape, self.hidden_layers, self.dropouts = tuple(feature_shape), list(hidden_layers), list(dropouts)
        self._record_config(feature_shape=list(feature_shape), hidden_layers=self.hidden_layers, dropouts=self.dropouts)
        self.model = self.build_model()

    def build_model(self):
        model = Sequential()
        model.add(InputLayer(input_shape=self.feature_shape))
        model.add(LSTM(self.hidden_layers[0], return_sequences=True))
        for i in range(1, len(self.hidden_layers) - 1):
            model.add(LSTM(self.hidden_layers[i], return_sequences=True)).add(Dropout(self.dropouts[i]))
        model.add(LSTM(self.hidden_layers[-1], return_sequences=False)).add(Dropout(self.dropouts[-1]))
        model.add(Dense(1))
        return model

    def forward(self, x):
        return self.model(x)

    @classmethod
    def unroll(cls, data_rdd, unroll_length, predict_step=1):
        """[n, features] array (or iterable of rows) -> list of FeatureLableIndex: window i covers rows
        [i, i+unroll_length), its label is column 0 of row i + unroll_length + predict_step - 1."""
        data = np.asarray(list(data_rdd) if not isinstance(data_rdd, np.ndarray) else data_rdd, dtype=np.float32)
        if data.ndim == 1:
            data = data[:, None]
        n = len(data) - unroll_length - predict_step + 1
        return [FeatureLableIndex(data[i:i + unroll_length], float(data[i + unroll_length + predict_step - 1, 0]), i) for i in range(max(n, 0))]

    @staticmethod
    def to_arrays(unrolled):
        return np.stack([u.feature for u in unrolled]), np.asarray([[u.label] for u in unrolled], dtype=np.float32)

    @classmethod
    def detect_anomalies(cls, ytruth, ypredict, anomaly_size):
        """-> list of (truth, prediction, anomaly-or-None) where the `anomaly_size` largest absolute errors are flagged."""
        yt, yp = np.asarray(ytruth, dtype=np.float64).reshape(-1), np.asarray(ypredict, dtype=np.float64).reshape(-1)
        if len(yt) != len(yp):
            raise ValueError("ytruth and ypredict must have the same length")
        diff = np.abs(yt - yp)
        k = min(int(anomaly_size), len(diff))
        threshold = np.sort(diff)[::-1][k - 1] if k > 0 else np.inf
        return [(float(t), float(p), float(t) if d >= threshold else None) for t, p, d in zip(yt, yp, diff)]

    @staticmethod
    def standardScale(df):
        return (df - df.mean()) / df.std()

    @staticmethod
    def train_test_split(unrolled, test_size):
        unrolled = list(unrolled)
        cut = len(unrolled) - int(test_size)
        return unrolled[:cut], unrolled[cut:]
