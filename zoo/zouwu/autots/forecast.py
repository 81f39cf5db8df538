"""AutoTSTrainer / TSPipeline (reference: pyzoo/zoo/zouwu/autots/forecast.py:22-168)."""
from __future__ import annotations

from ...automl.config.recipe import *  # noqa: F401,F403  (recipes are re-exported here, as in the reference)
from ...automl.config.recipe import Recipe, SmokeRecipe
from ...automl.pipeline.time_sequence import load_ts_pipeline
from ...automl.regression.time_sequence_predictor import TimeSequencePredictor


class AutoTSTrainer:
    """The automated time-series forecast trainer."""

    def __init__(self, horizon=1, dt_col="datetime", target_col="value", extra_features_col=None):
        self.internal = TimeSequencePredictor(dt_col=dt_col, target_col=target_col, future_seq_len=horizon,
                                              extra_features_col=extra_features_col)

    def fit(self, train_df, validation_df=None, metric="mse", recipe: Recipe = None, uncertainty: bool = False,
            distributed: bool = False, hdfs_url=None):
        zoo_pipeline = self.internal.fit(train_df, validation_df, metric, recipe or SmokeRecipe(), mc=uncertainty,
                                         distributed=distributed, hdfs_url=hdfs_url)
        ppl = TSPipeline()
        ppl.internal = zoo_pipeline
        ppl.uncertainty = uncertainty
        return ppl


class TSPipeline:
    """A pipeline for time-series forecasting: obtained from `AutoTSTrainer.fit` or `TSPipeline.load`."""

    def __init__(self):
        self.internal = None
        self.uncertainty = False

    def save(self, pipeline_file):
        return self.internal.save(pipeline_file)

    @staticmethod
    def load(pipeline_file):
        tsppl = TSPipeline()
        tsppl.internal = load_ts_pipeline(pipeline_file)
        return tsppl

    def fit(self, input_df, validation_df=None, uncertainty: bool = False, epochs=1, **user_config):
        """Incremental fitting; with `user_config` the pipeline is re-trained from scratch with those settings."""
        if user_config:
            return self.internal.fit_with_fixed_configs(input_df=input_df, validation_df=validation_df, mc=uncertainty,
                                                        epochs=epochs, **user_config)
        return self.internal.fit(input_df=input_df, validation_df=validation_df, mc=uncertainty, epoch_num=epochs)

    def predict(self, input_df):
        return self.internal.predict(input_df)

    def predict_with_uncertainty(self, input_df, n_iter=100):
        return self.internal.predict_with_uncertainty(input_df, n_iter)

    def evaluate(self, input_df, metrics=("mse",), multioutput="raw_values"):
        return self.internal.evaluate(input_df, list(metrics), multioutput=multioutput)

    def describe(self):
        return self.internal.describe()
