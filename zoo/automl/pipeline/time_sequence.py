"""TimeSequencePipeline: feature transformer + model + config as one object
(reference: pyzoo/zoo/automl/pipeline/time_sequence.py:28-221): `fit` (incremental), `fit_with_fixed_configs`,
`evaluate`, `predict`, `predict_with_uncertainty`, `save` / `load_ts_pipeline` (a zip of model file + config JSON)."""
from __future__ import annotations

import time

import numpy as np

from ..common.metrics import Evaluator
from ..common.util import restore_zip, save_zip
from ..feature.time_sequence import TimeSequenceFeatureTransformer
from ..model.time_sequence import TimeSequenceModel


class TimeSequencePipeline:
    def __init__(self, feature_transformers=None, model=None, config=None, name=None):
        self.feature_transformers, self.model, self.config = feature_transformers, model, config or {}
        self.name = name
        self.time = time.strftime("%Y%m%d-%H%M%S")

    def describe(self):
        init_info = ["future_seq_len", "dt_col", "target_col", "extra_features_col", "drop_missing"]
        print("**** Initialization info ****")
        for k in init_info:
            print(k + ":", getattr(self.feature_transformers, k))
        print("**** Training config ****")
        print(self.config)

    def fit(self, input_df, validation_df=None, mc=False, epoch_num=20):
        """Continue training the pipeline's model on new data (the scaler is re-fitted, as in the reference)."""
        x, y = self.feature_transformers.fit_transform(input_df, **self.config)
        val = self.feature_transformers.transform(validation_df, is_train=True) if validation_df is not None else None
        cfg = dict(self.config, epochs=1)
        new_metric = None
        for i in range(epoch_num):
            new_metric = self.model.fit_eval(x, y, validation_data=val, mc=mc, verbose=1, **cfg)
            print(f"Epoch {i + 1}/{epoch_num}: {cfg.get('metric', 'mse')} {new_metric:.6f}")
        print("Fit done!")
        return new_metric

    def fit_with_fixed_configs(self, input_df, validation_df=None, mc=False, **user_configs):
        """Train from scratch with the given config instead of searching (time_sequence.py:100-136)."""
        self.config = dict(self.config, **user_configs)
        ft = self.feature_transformers
        if "selected_features" not in self.config:
            self.config["selected_features"] = ft.get_feature_list(input_df)
        x, y = ft.fit_transform(input_df, **self.config)
        val = ft.transform(validation_df, is_train=True) if validation_df is not None else None
        self.model = TimeSequenceModel(check_optional_config=False, future_seq_len=ft.future_seq_len)
        return self.model.fit_eval(x, y, validation_data=val, mc=mc, **self.config)

    def evaluate(self, input_df, metrics=("mse",), multioutput="raw_values"):
        if isinstance(metrics, str):
            metrics = [metrics]
        for m in metrics:
            Evaluator.check_metric(m)
        x, _ = self.feature_transformers.transform(input_df, is_train=True)
        y_pred = self.model.predict(x)
        y_unscale, y_true = self.feature_transformers.post_processing(input_df, y_pred, is_train=True)
        return [Evaluator.evaluate(m, y_true, y_unscale, multioutput=multioutput) for m in metrics]

    def predict(self, input_df):
        x, _ = self.feature_transformers.transform(input_df, is_train=False)
        return self.feature_transformers.post_processing(input_df, self.model.predict(x), is_train=False)

    def predict_with_uncertainty(self, input_df, n_iter=100):
        x, _ = self.feature_transformers.transform(input_df, is_train=False)
        mean, std = self.model.predict_with_uncertainty(x, n_iter)
        out = self.feature_transformers.post_processing(input_df, mean, is_train=False)
        return out, self.feature_transformers.unscale_uncertainty(std)

    def save(self, ppl_file=None):
        ppl_file = ppl_file or f"{self.name or 'zoo_ts_pipeline'}_{self.time}.ppl"
        save_zip(ppl_file, self.feature_transformers, self.model, self.config)
        print("Pipeline is saved in", ppl_file)
        return ppl_file


def load_ts_pipeline(file):
    ft = TimeSequenceFeatureTransformer()
    model = TimeSequenceModel(check_optional_config=False)
    all_config = restore_zip(file, ft, model)
    ppl = TimeSequencePipeline(feature_transformers=ft, model=model, config=all_config)
    print("Restore pipeline from", file)
    return ppl
