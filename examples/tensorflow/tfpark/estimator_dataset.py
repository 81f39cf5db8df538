"""TFEstimator.from_model_fn (reference example: pyzoo/zoo/examples/tensorflow/tfpark/estimator/estimator_dataset.py: a model_fn
building LeNet logits + sparse softmax loss, `ZooOptimizer(tf.train.AdamOptimizer())` as the train op, `estimator.train(input_fn,
steps=...)`, `estimator.evaluate(input_fn, ["acc"])`, `estimator.predict(input_fn)`).  model_fn returns a TFEstimatorSpec holding
a torch module; state lives in model_dir between the three calls."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.pipeline.api.keras.optimizers import Adam
from zoo.tfpark import TFDataset, TFEstimator, TFEstimatorSpec, ZooOptimizer

torch.manual_seed(0)
rng = np.random.RandomState(1)
x = rng.randn(3000, 20).astype(np.float32)
y = (x[:, :10].sum(1) > x[:, 10:].sum(1)).astype(np.int64)


def model_fn(features, labels, mode, params):
    net = nn.Sequential(nn.Linear(features.shape[1], params["hidden"]), nn.ReLU(), nn.Linear(params["hidden"], 2))
    return TFEstimatorSpec(mode, model=net, loss=nn.CrossEntropyLoss(), train_op=ZooOptimizer(Adam(lr=5e-3)),
                           predictions=lambda logits: {"classes": logits.argmax(-1), "probabilities": logits.softmax(-1)})


def input_fn(mode):
    if mode == "train":
        return TFDataset.from_ndarrays((x[:2400], y[:2400]), batch_size=120)
    if mode == "eval":
        return TFDataset.from_ndarrays((x[2400:], y[2400:]), batch_per_thread=200)
    return TFDataset.from_ndarrays(x[2400:], batch_per_thread=200)


estimator = TFEstimator.from_model_fn(model_fn, model_dir=tempfile.mkdtemp(), params={"hidden": 32})
estimator.train(input_fn, steps=int(os.environ.get("STEPS", 200)))
print("evaluation:", estimator.evaluate(input_fn, ["acc", "loss"]))
out = estimator.predict(input_fn, predict_keys=["classes"])
print("predicted %d records at global step %d; first classes %s" % (len(out), estimator.global_step, [int(o["classes"]) for o in out[:8]]))
