"""NNFrames transfer learning (reference examples: pyzoo/zoo/examples/nnframes/imageTransferLearning/
ImageTransferLearningExample.py and finetune/image_finetuning_example.py: a pre-trained backbone as a fixed feature extractor
(`NNModel(...).transform` adds an embedding column), then `NNClassifier` trains a small head on the embeddings; the two stages form
one pipeline)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import pandas as pd
import torch
import torch.nn as nn

from zoo.feature.common import ChainedPreprocessing, SeqToTensor
from zoo.feature.image import ImageChannelNormalize, ImageFeatureToTensor, ImageMatToTensor, ImageResize, RowToImageFeature
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.api.keras.optimizers import Adam
from zoo.pipeline.nnframes import NNClassifier, NNModel

rng = np.random.RandomState(0)
rows = []
for i in range(int(os.environ.get("IMAGES", 96))):
    label = i % 2
    img = rng.randint(0, 80, (48, 48, 3), dtype=np.uint8)
    img[8:40, 8:40, label * 2] += 120
    rows.append({"image": {"origin": f"mem:{i}", "height": 48, "width": 48, "nChannels": 3, "mode": 16, "data": img}, "label": float(label + 1)})
df = pd.DataFrame(rows)
train, val = df.iloc[: len(df) * 3 // 4], df.iloc[len(df) * 3 // 4:]

torch.manual_seed(0)
backbone = nn.Sequential(nn.Conv2d(3, 16, 5, stride=2), nn.ReLU(), nn.Conv2d(16, 32, 3, stride=2), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten()).eval()
chain = ChainedPreprocessing([RowToImageFeature(), ImageResize(40, 40), ImageChannelNormalize(100.0, 100.0, 100.0, 60.0, 60.0, 60.0),
                              ImageMatToTensor(), ImageFeatureToTensor()])
featurizer = NNModel(backbone, chain).setFeaturesCol("image").setPredictionCol("embedding").setBatchSize(32)     # stage 1: frozen extractor
head = nn.Sequential(nn.Linear(32, 2), nn.LogSoftmax(-1))
classifier = NNClassifier(head, SparseCategoricalCrossEntropy(log_prob_as_input=True, zero_based_label=False),
        SeqToTensor([32])).setFeaturesCol("embedding").setBatchSize(24) \
    .setMaxEpoch(int(os.environ.get("EPOCHS", 30))).setOptimMethod(Adam(lr=5e-2))                                   # stage 2: trainable head
model = classifier.fit(featurizer.transform(train))
pred = model.transform(featurizer.transform(val))
print("validation accuracy: %.3f over %d images" % (float((pred["prediction"] == pred["label"]).mean()), len(pred)))
