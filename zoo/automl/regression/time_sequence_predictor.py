"""TimeSequencePredictor: hyper-parameter search that returns the best TimeSequencePipeline
(reference: pyzoo/zoo/automl/regression/time_sequence_predictor.py:37-296)."""
from __future__ import annotations

import pandas as pd

from ..common.metrics import Evaluator
from ..common.util import restore_zip
from ..config.recipe import SmokeRecipe
from ..feature.time_sequence import TimeSequenceFeatureTransformer
from ..model.time_sequence import TimeSequenceModel
from ..pipeline.time_sequence import TimeSequencePipeline
from ..search.RayTuneSearchEngine import RayTuneSearchEngine


class TimeSequencePredictor:
    def __init__(self, name="automl", logs_dir="~/zoo_automl_logs", future_seq_len=1, dt_col="datetime", target_col="value",
                 extra_features_col=None, drop_missing=True):
        import os

        self.name, self.logs_dir = name, os.path.expanduser(logs_dir)
        self.future_seq_len, self.dt_col, self.target_col = future_seq_len, dt_col, target_col
        self.extra_features_col, self.drop_missing = extra_features_col, drop_missing
        self.pipeline = None

    def _check_df(self, df):
        if isinstance(df, list):
            for d in df:
                self._check_df(d)
            return
        if not isinstance(df, pd.DataFrame):
            raise ValueError("input_df should be a pandas DataFrame (or a list of them)")
        for c in [self.dt_col, self.target_col] + list(self.extra_features_col or []):
            if c not in df.columns:
                raise ValueError(f"column {c} is not in the input data")

    def fit(self, input_df, validation_df=None, metric="mse", recipe=SmokeRecipe(), mc=False, resources_per_trial=None,
            distributed=False, hdfs_url=None):
        """Search with `recipe`; returns the TimeSequencePipeline of the best trial."""
        self._check_df(input_df)
        if validation_df is not None:
            self._check_df(validation_df)
        Evaluator.check_metric(metric)
        self.pipeline = self._hp_search(input_df, validation_df, metric, recipe, mc, resources_per_trial or {"cpu": 2})
        return self.pipeline

    def evaluate(self, input_df, metric=None):
        if self.pipeline is None:
            raise RuntimeError("call fit first")
        return self.pipeline.evaluate(input_df, metric or ["mse"])

    def predict(self, input_df):
        if self.pipeline is None:
            raise RuntimeError("call fit first")
        return self.pipeline.predict(input_df)

    def _hp_search(self, input_df, validation_df, metric, recipe, mc, resources_per_trial):
        ft = TimeSequenceFeatureTransformer(self.future_seq_len, self.dt_col, self.target_col, self.extra_features_col, self.drop_missing)
        feature_list = ft.get_feature_list(input_df if not isinstance(input_df, list) else input_df[0])
        fsl = self.future_seq_len

        def model_create():
            return TimeSequenceModel(check_optional_config=False, future_seq_len=fsl)

        search_space = recipe.search_space(feature_list)
        engine = RayTuneSearchEngine(logs_dir=self.logs_dir, resources_per_trial=resources_per_trial, name=self.name)
        engine.compile(input_df, model_create, search_space, recipe=recipe, feature_transformers=ft, future_seq_len=fsl,
                       validation_df=validation_df, mc=mc, metric=metric)
        engine.run()
        best = engine.get_best_trials(k=1)
        if not best:
            raise RuntimeError("hyper-parameter search produced no finished trial")
        return self._make_pipeline(best[0], ft)

    def _make_pipeline(self, trial, ft):
        model = TimeSequenceModel(check_optional_config=False, future_seq_len=self.future_seq_len)
        all_config = restore_zip(trial.model_path, ft, model)
        return TimeSequencePipeline(name=self.name, feature_transformers=ft, model=model, config=all_config)
