"""AutoTS: search a forecasting pipeline on an hourly series, then predict / evaluate / save / load it."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import pandas as pd

from zoo.automl.config.recipe import LSTMGridRandomRecipe, TCNSmokeRecipe
from zoo.zouwu.autots import AutoTSTrainer, TSPipeline

ap = argparse.ArgumentParser()
ap.add_argument("--recipe", default="lstm", choices=["lstm", "tcn"])
ap.add_argument("--points", type=int, default=1200)
a = ap.parse_args()
t = np.arange(a.points)
df = pd.DataFrame({"datetime": pd.date_range("2019-01-01", periods=a.points, freq="h"),
                   "value": 10 + 3 * np.sin(2 * np.pi * t / 24) + 0.3 * np.random.RandomState(0).randn(a.points)})
train, val, test = df[:-300], df[-300:-100], df[-100:]
recipe = LSTMGridRandomRecipe(num_rand_samples=1, epochs=2, training_iteration=3, look_back=24, lstm_1_units=[16, 32], lstm_2_units=[16], batch_size=[64]) \
    if a.recipe == "lstm" else TCNSmokeRecipe()
ppl = AutoTSTrainer(horizon=1, dt_col="datetime", target_col="value").fit(train, val, metric="mse", recipe=recipe)
print("test mse / smape:", ppl.evaluate(test, metrics=["mse", "smape"]))
print(ppl.predict(test).head())
path = ppl.save("/tmp/zoo_autots_example.ppl")
print("reloaded prediction equal:", np.allclose(TSPipeline.load(path).predict(test)["value"], ppl.predict(test)["value"]))
