"""TimeSequenceFeatureTransformer (reference: pyzoo/zoo/automl/feature/time_sequence.py:30-573).

dataframe (datetime, target, extra features) -> calendar features (the reference derives them with
featuretools: MONTH / WEEKDAY / DAY / HOUR / IS_WEEKEND / IS_AWAKE / IS_BUSY_HOURS of the datetime column;
the same names are kept) -> standard scaling -> rolling windows `x [n, past_seq_len, 1+features]`,
`y [n, future_seq_len]`; `post_processing` un-scales predictions and attaches the forecast timestamps;
`save` / `restore` round-trip the scaler and the configuration through JSON.
"""
from __future__ import annotations

import numpy as np
import pandas as pd

from ..common.util import load_config, save_config


class TimeSequenceFeatureTransformer:
    def __init__(self, future_seq_len=1, dt_col="datetime", target_col="value", extra_features_col=None, drop_missing=True):
        self.future_seq_len, self.dt_col, self.target_col = int(future_seq_len), dt_col, target_col
        self.extra_features_col = list(extra_features_col) if extra_features_col else None
        self.drop_missing = drop_missing
        self.past_seq_len = None
        self.feature_data = None
        self.scale_mean = self.scale_std = None
        self.config = None

    # ---- features ------------------------------------------------------------------------------
    def _calendar(self, dt: pd.Series):
        c = self.dt_col
        hour = dt.dt.hour
        return {f"MONTH({c})": dt.dt.month, f"WEEKDAY({c})": dt.dt.weekday, f"DAY({c})": dt.dt.day, f"HOUR({c})": hour,
                f"IS_WEEKEND({c})": (dt.dt.weekday >= 5).astype(int),
                f"IS_AWAKE({c})": (((hour >= 6) & (hour <= 23)) | (hour == 0)).astype(int),
                f"IS_BUSY_HOURS({c})": (((hour >= 7) & (hour <= 9)) | ((hour >= 16) & (hour <= 19))).astype(int)}

    def get_feature_list(self, input_df=None):
        c = self.dt_col
        feats = [f"MONTH({c})", f"WEEKDAY({c})", f"DAY({c})", f"HOUR({c})", f"IS_WEEKEND({c})", f"IS_AWAKE({c})", f"IS_BUSY_HOURS({c})"]
        return feats + (self.extra_features_col or [])

    def _check_input(self, input_df, mode="train"):
        df = input_df.copy()
        if not isinstance(df, pd.DataFrame):
            raise ValueError("input should be a pandas DataFrame")
        for col in [self.dt_col, self.target_col] + (self.extra_features_col or []):
            if col not in df.columns:
                raise ValueError(f"column {col} is not in the input data")
        if not pd.api.types.is_datetime64_any_dtype(df[self.dt_col]):
            raise ValueError(f"the dtype of datetime column is required to be datetime64, got {df[self.dt_col].dtype}")
        if df[self.dt_col].isnull().any():
            raise ValueError("missing datetime in input data")
        if df[self.target_col].isnull().any() or (self.extra_features_col and df[self.extra_features_col].isnull().values.any()):
            if not self.drop_missing:
                raise ValueError("missing values in input data")
            df = df.dropna(subset=[self.target_col] + (self.extra_features_col or []))
        min_len = (self.past_seq_len or 1) + (self.future_seq_len if mode == "train" else 0)
        if len(df) < min_len:
            raise ValueError(f"Length of {mode} data should be at least {min_len} (past_seq_len + future_seq_len), got {len(df)}")
        return df.reset_index(drop=True)

    def _features(self, df, selected):
        cal = self._calendar(df[self.dt_col])
        cols = {self.target_col: df[self.target_col].astype(float)}
        for f in selected:
            cols[f] = (cal[f] if f in cal else df[f]).astype(float)
        return pd.DataFrame(cols)

    def _get_feat_config(self, **config):
        if "selected_features" not in config:
            raise ValueError("selected_features is required in the config")
        sel = config["selected_features"]
        if isinstance(sel, str):
            import json

            sel = json.loads(sel)
        self.past_seq_len = int(config.get("past_seq_len", 2))
        return {"selected_features": list(sel), "past_seq_len": self.past_seq_len}

    # ---- rolling ---------------------------------------------------------------------------------
    @staticmethod
    def _roll(data, seq_len):
        n = data.shape[0] - seq_len + 1
        idx = np.arange(seq_len)[None, :] + np.arange(n)[:, None]
        return data[idx]

    def _roll_train(self, data):
        p, f = self.past_seq_len, self.future_seq_len
        x = self._roll(data[:-f], p)
        y = self._roll(data[p:, 0], f)
        return x.astype(np.float32), y.astype(np.float32)

    def _roll_test(self, data):
        return self._roll(data, self.past_seq_len).astype(np.float32)

    # ---- API ---------------------------------------------------------------------------------------
    def _fit_transform(self, input_df):
        df = self._check_input(input_df, "train")
        feat = self._features(df, self.config["selected_features"]).values
        self.scale_mean, self.scale_std = feat.mean(0), np.where(feat.std(0) == 0, 1.0, feat.std(0))
        return self._roll_train((feat - self.scale_mean) / self.scale_std)

    def fit_transform(self, input_df, **config):
        self.config = self._get_feat_config(**config)
        if isinstance(input_df, list):
            # fit the scaler on all frames together, then window each frame separately
            dfs = [self._check_input(d, "train") for d in input_df]
            feats = [self._features(d, self.config["selected_features"]).values for d in dfs]
            allf = np.concatenate(feats)
            self.scale_mean, self.scale_std = allf.mean(0), np.where(allf.std(0) == 0, 1.0, allf.std(0))
            xs, ys = zip(*[self._roll_train((f - self.scale_mean) / self.scale_std) for f in feats])
            return np.concatenate(xs), np.concatenate(ys)
        return self._fit_transform(input_df)

    def transform(self, input_df, is_train=True):
        if self.config is None or self.scale_mean is None:
            raise Exception("Needs to call fit_transform or restore first before calling transform")

        def one(df):
            df = self._check_input(df, "train" if is_train else "test")
            feat = (self._features(df, self.config["selected_features"]).values - self.scale_mean) / self.scale_std
            return self._roll_train(feat) if is_train else (self._roll_test(feat), None)

        if isinstance(input_df, list):
            xs, ys = zip(*[one(d) for d in input_df])
            return np.concatenate(xs), (np.concatenate(ys) if is_train else None)
        return one(input_df)

    def _unscale(self, y):
        return y * self.scale_std[0] + self.scale_mean[0]

    def unscale_uncertainty(self, y_uncertainty):
        return y_uncertainty * self.scale_std[0]

    def _get_y_pred_dt(self, input_df):
        """Timestamps of the forecasts: one row per test window, `future_seq_len` steps after its last input."""
        df = self._check_input(input_df, "test")
        dt = df[self.dt_col].values
        step = dt[-1] - dt[-2] if len(dt) > 1 else np.timedelta64(1, "D")
        last = dt[self.past_seq_len - 1:]
        return last, step

    def post_processing(self, input_df, y_pred, is_train):
        """Train: un-scaled (y_pred, y_true) arrays.  Test: DataFrame(datetime, value[_k])."""
        y_unscale = self._unscale(np.asarray(y_pred))
        if is_train:
            def truth(df):
                _, y = self.transform(df, is_train=True)
                return self._unscale(y)

            y_true = np.concatenate([truth(d) for d in input_df]) if isinstance(input_df, list) else truth(input_df)
            return y_unscale.reshape(y_true.shape), y_true

        def frame(df, yp):
            last, step = self._get_y_pred_dt(df)
            yp = yp.reshape(len(last), -1)
            out = pd.DataFrame({self.dt_col: last + step})
            if yp.shape[1] == 1:
                out[self.target_col] = yp[:, 0]
            else:
                for k in range(yp.shape[1]):
                    out[f"{self.target_col}_{k}"] = yp[:, k]
            return out

        if isinstance(input_df, list):
            outs, s = [], 0
            for d in input_df:
                n = len(self._check_input(d, "test")) - self.past_seq_len + 1
                outs.append(frame(d, y_unscale[s:s + n]))
                s += n
            return outs
        return frame(input_df, y_unscale)

    def save(self, file_path, replace=False):
        data = {"mean": self.scale_mean.tolist(), "scale": self.scale_std.tolist(), "future_seq_len": self.future_seq_len,
                "dt_col": self.dt_col, "target_col": self.target_col, "extra_features_col": self.extra_features_col,
                "drop_missing": self.drop_missing}
        data.update(self.config or {})
        save_config(file_path, data, replace=replace)

    def restore(self, **config):
        self.scale_mean, self.scale_std = np.asarray(config["mean"], dtype=np.float64), np.asarray(config["scale"], dtype=np.float64)
        self.future_seq_len, self.dt_col, self.target_col = config["future_seq_len"], config["dt_col"], config["target_col"]
        self.extra_features_col, self.drop_missing = config.get("extra_features_col"), config.get("drop_missing", True)
        self.config = self._get_feat_config(**config)
        return self

    @classmethod
    def load(cls, file_path):
        return cls().restore(**load_config(file_path))
