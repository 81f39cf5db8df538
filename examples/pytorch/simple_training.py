"""The smallest TorchNet + TorchCriterion + NNClassifier training run (reference example:
pyzoo/zoo/examples/pytorch/train/SimpleTrainingExample.py: a 2-4-1 torch model, MSE criterion, four rows of a DataFrame, Adam,
`classifier.fit(df)` then `transform(df)`)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import pandas as pd
import torch
import torch.nn as nn

from zoo.common.nncontext import init_nncontext
from zoo.pipeline.api.keras.optimizers import Adam
from zoo.pipeline.api.net import TorchCriterion, TorchNet
from zoo.pipeline.nnframes import NNClassifier


class SimpleTorchModel(nn.Module):
    def __init__(self):
        super().__init__()
        self.dense1 = nn.Linear(2, 4)
        self.dense2 = nn.Linear(4, 1)

    def forward(self, x):
        return torch.sigmoid(self.dense2(self.dense1(x)))


if __name__ == "__main__":
    init_nncontext("testNNClassifer")
    torch.manual_seed(0)
    df = pd.DataFrame({"features": [[2.0, 1.0], [1.0, 2.0], [2.0, 1.0], [1.0, 2.0]], "label": [1.0, 0.0, 1.0, 0.0]})
    az_model = TorchNet.from_pytorch(SimpleTorchModel(), [1, 2])
    az_criterion = TorchCriterion.from_pytorch(nn.MSELoss(), [1, 1], [1, 1])
    classifier = NNClassifier(az_model, az_criterion).setBatchSize(4).setOptimMethod(Adam()).setLearningRate(0.05) \
        .setMaxEpoch(int(os.environ.get("SIMPLE_EPOCHS", 200)))
    nn_classifier_model = classifier.fit(df)
    print("After training: ")
    res = nn_classifier_model.transform(df)
    print(res.to_string())
    print("predictions match labels:", [float(p) for p in res["prediction"]] == list(df["label"]))
