"""Wide&Deep on Census-shaped columns with per-sub-module optimizers (FTRL for the wide part, Adagrad for the deep part)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch

from zoo.common.nncontext import init_nncontext
from zoo.common.triggers import MaxEpoch
from zoo.feature.common import FeatureSet
from zoo.models.recommendation import ColumnFeatureInfo, WideAndDeep
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.api.keras.optimizers import Adagrad, Ftrl
from zoo.pipeline.estimator import Estimator

ap = argparse.ArgumentParser()
ap.add_argument("--epochs", type=int, default=2)
ap.add_argument("--batch-size", type=int, default=2048)
ap.add_argument("--rows", type=int, default=32561)
a = ap.parse_args()
init_nncontext("WideAndDeep example")
info = ColumnFeatureInfo(wide_base_cols=["edu", "mari", "rela", "work", "occ", "age_bucket"], wide_base_dims=[16, 7, 6, 9, 1000, 11],
                         wide_cross_cols=["edu_occ", "age_edu_occ"], wide_cross_dims=[1000, 1000], indicator_cols=["work", "edu", "mari", "rela"],
                         indicator_dims=[9, 16, 7, 6], embed_cols=["occ"], embed_in_dims=[1000], embed_out_dims=[8],
                         continuous_cols=["age", "education_num", "capital_gain", "capital_loss", "hours_per_week"])
model = WideAndDeep(2, info, model_type="wide_n_deep", hidden_layers=[100, 75, 50, 25], sparse=False)
rng = np.random.RandomState(0)
xs = model.example_inputs(a.rows, rng)
y = torch.from_numpy(((xs[0][:, 0] + xs[2][:, 0]) % 2 + 1).numpy().astype(np.int64))      # 1-based labels
est = Estimator(model, {"wide_part": Ftrl(learningrate=5e-3), "deep_part": Adagrad(learningrate=1e-2)})
train = FeatureSet(features=list(xs), labels=y)
est.train(train, SparseCategoricalCrossEntropy(zero_based_label=False), MaxEpoch(a.epochs), batch_size=a.batch_size)
print(est.evaluate(train, ["accuracy", "loss"], batch_size=a.batch_size))
