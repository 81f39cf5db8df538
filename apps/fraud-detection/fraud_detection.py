"""Fraud detection on an imbalanced transaction table (reference app: apps/fraud-detection/fraud-detection.ipynb -- credit-card
data, bagging of NNClassifiers trained on under-sampled splits in a Spark ML pipeline, AUPRC / recall on the held-out set).
Synthetic table with 0.5 % positives; DataFrame in, DataFrame out, through NNClassifier."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import pandas as pd
import torch.nn as nn

from zoo.pipeline.api.keras.optimizers import Adam
from zoo.pipeline.api.net import TorchCriterion
from sklearn.metrics import average_precision_score, recall_score

from zoo.pipeline.nnframes import NNClassifier

rng = np.random.RandomState(0)
n, d = 40000, 16
x = rng.randn(n, d).astype(np.float32)
fraud = rng.rand(n) < 0.005
x[fraud, :4] += 2.5                                            # frauds sit in a corner of feature space
df = pd.DataFrame({"features": list(x), "label": fraud.astype(np.float64) + 1.0})     # labels are 1-based like the reference
train, test = df.iloc[:32000], df.iloc[32000:]

models = []
pos = train[train.label == 2.0]
for bag in range(int(os.environ.get("BAGS", 3))):             # bagging over under-sampled negatives
    neg = train[train.label == 1.0].sample(len(pos) * 10, random_state=bag)
    part = pd.concat([pos, neg]).sample(frac=1.0, random_state=bag)
    net = nn.Sequential(nn.Linear(d, 32), nn.ReLU(), nn.Linear(32, 2))
    crit = TorchCriterion.from_pytorch(lambda out, y: nn.functional.cross_entropy(out, (y.reshape(-1) - 1).long()))
    clf = NNClassifier(net, crit, [d]).setBatchSize(256).setMaxEpoch(int(os.environ.get("EPOCHS", 8))).setOptimMethod(Adam(lr=3e-3))
    models.append(clf.fit(part))
votes = np.mean([np.asarray(m.transform(test)["prediction"].tolist()) == 2.0 for m in models], axis=0)
pred = votes >= 0.5
truth = (test.label == 2.0).to_numpy()
print("held-out frauds %d, recall %.3f, precision %.3f, AUPRC(votes) %.3f" % (truth.sum(), recall_score(truth, pred),
      (pred & truth).sum() / max(pred.sum(), 1), average_precision_score(truth, votes)))
