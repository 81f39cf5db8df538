"""Anomaly detection on a univariate series (NAB NYC-taxi shaped): unroll -> stacked LSTM regressor -> top-N errors."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.models.anomalydetection import AnomalyDetector

t = np.arange(3000, dtype=np.float32)
series = np.sin(t / 24 * 2 * np.pi) + 0.05 * np.random.RandomState(0).randn(3000).astype(np.float32)
series[2800] += 3.0                                                    # an injected anomaly
scaled = (series - series.mean()) / series.std()
unrolled = AnomalyDetector.unroll(scaled[:, None], unroll_length=24)
train, test = AnomalyDetector.train_test_split(unrolled, 500)
x, y = AnomalyDetector.to_arrays(train)
model = AnomalyDetector(feature_shape=(24, 1), hidden_layers=[8, 32, 15], dropouts=[0.2, 0.2, 0.2])
model.compile(optimizer="adam", loss="mse")
model.fit(x, y, batch_size=256, nb_epoch=3)
xt, yt = AnomalyDetector.to_arrays(test)
anomalies = AnomalyDetector.detect_anomalies(yt, model.predict(xt), anomaly_size=3)
print("flagged test positions:", [i for i, a in enumerate(anomalies) if a[2] is not None], "(injected at", 2800 - 24 - len(train), ")")
