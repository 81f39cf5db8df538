"""Evaluation metrics of the AutoML time-series stack (reference: pyzoo/zoo/automl/common/metrics.py:24-245):
ME, MAE, MSE, RMSE, MSLE, R2, MPE, MAPE, MSPE, sMAPE, MDAPE, sMDAPE -- with the reference's `multioutput`
('raw_values' | 'uniform_average') convention."""
from __future__ import annotations

import numpy as np

EPSILON = 1e-10


def _prep(y_true, y_pred):
    y_true, y_pred = np.asarray(y_true, dtype=np.float64), np.asarray(y_pred, dtype=np.float64)
    if y_true.shape != y_pred.shape:
        if y_true.size == y_pred.size:
            y_pred = y_pred.reshape(y_true.shape)
        else:
            raise ValueError(f"y_true and y_pred have different shapes: {y_true.shape} vs {y_pred.shape}")
    if y_true.ndim == 1:
        y_true, y_pred = y_true[:, None], y_pred[:, None]
    return y_true, y_pred


def _out(v, multioutput):
    if multioutput == "raw_values":
        return v
    if multioutput == "uniform_average":
        return float(np.mean(v))
    raise ValueError("multioutput must be 'raw_values' or 'uniform_average'")


def _metric(fn):
    def wrapped(y_true, y_pred, multioutput="raw_values"):
        t, p = _prep(y_true, y_pred)
        return _out(fn(t, p), multioutput)

    wrapped.__name__ = fn.__name__
    return wrapped


ME = _metric(lambda t, p: np.mean(t - p, axis=0))
MAE = _metric(lambda t, p: np.mean(np.abs(t - p), axis=0))
MSE = _metric(lambda t, p: np.mean((t - p) ** 2, axis=0))
RMSE = _metric(lambda t, p: np.sqrt(np.mean((t - p) ** 2, axis=0)))
MSLE = _metric(lambda t, p: np.mean((np.log1p(t) - np.log1p(p)) ** 2, axis=0))
MPE = _metric(lambda t, p: 100 * np.mean((t - p) / (t + EPSILON), axis=0))
MAPE = _metric(lambda t, p: 100 * np.mean(np.abs((t - p) / (t + EPSILON)), axis=0))
MSPE = _metric(lambda t, p: 100 * np.mean(((t - p) / (t + EPSILON)) ** 2, axis=0))
sMAPE = _metric(lambda t, p: 100 * np.mean(np.abs(t - p) / (np.abs(t) + np.abs(p) + EPSILON), axis=0))
MDAPE = _metric(lambda t, p: 100 * np.median(np.abs((t - p) / (t + EPSILON)), axis=0))
sMDAPE = _metric(lambda t, p: 100 * np.median(np.abs(t - p) / (np.abs(t) + np.abs(p) + EPSILON), axis=0))


def _r2(t, p):
    ss_res = np.sum((t - p) ** 2, axis=0)
    ss_tot = np.sum((t - t.mean(axis=0)) ** 2, axis=0)
    return 1.0 - ss_res / np.where(ss_tot == 0, 1.0, ss_tot)


R2 = _metric(_r2)


class Evaluator:
    metrics_func = {"me": ME, "mae": MAE, "mse": MSE, "rmse": RMSE, "msle": MSLE, "r2": R2, "mpe": MPE, "mape": MAPE, "mspe": MSPE,
                    "smape": sMAPE, "mdape": MDAPE, "smdape": sMDAPE, "mean_squared_error": MSE, "mean_absolute_error": MAE,
                    "r_square": R2}
    max_mode = {"r2", "r_square"}

    @staticmethod
    def check_metric(metric):
        if metric.lower() not in Evaluator.metrics_func:
            raise ValueError(f"metric {metric} is not supported: choose from {sorted(Evaluator.metrics_func)}")

    @staticmethod
    def evaluate(metric, y_true, y_pred, multioutput="raw_values"):
        Evaluator.check_metric(metric)
        return Evaluator.metrics_func[metric.lower()](y_true, y_pred, multioutput=multioutput)

    @staticmethod
    def get_metric_mode(metric):
        return "max" if metric.lower() in Evaluator.max_mode else "min"


def check_input(y_true, y_pred, multioutput="raw_values"):
    t, p = _prep(y_true, y_pred)
    if multioutput not in ("raw_values", "uniform_average"):
        raise ValueError("multioutput must be 'raw_values' or 'uniform_average'")
    return t, p, t[0].shape
