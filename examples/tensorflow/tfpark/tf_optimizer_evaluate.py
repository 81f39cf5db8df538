"""Evaluate a trained model with TFPredictor over a TFDataset (reference example:
pyzoo/zoo/examples/tensorflow/tfpark/tf_optimizer/evaluate.py: restore the LeNet checkpoint written by train.py, build a
`TFDataset.from_rdd(..., batch_per_thread=20)`, `TFPredictor.from_outputs(sess, [correct])`, average the per-record hits).
Self-contained: trains for one short pass first and evaluates the checkpoint it wrote."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.common.triggers import MaxEpoch
from zoo.pipeline.api.keras.optimizers import Adam
from zoo.tfpark import TFDataset, TFOptimizer, TFPredictor

torch.manual_seed(0)
rng = np.random.RandomState(0)
N, CLASSES = int(os.environ.get("ROWS", 2048)), 10
protos = rng.rand(CLASSES, 1, 28, 28).astype(np.float32)
labels = rng.randint(0, CLASSES, N).astype(np.int64)
images = protos[labels] + 0.3 * rng.randn(N, 1, 28, 28).astype(np.float32)
split = N * 3 // 4


def lenet():
    return nn.Sequential(nn.Conv2d(1, 6, 5), nn.Tanh(), nn.MaxPool2d(2), nn.Conv2d(6, 12, 5), nn.Tanh(), nn.MaxPool2d(2), nn.Flatten(),
                         nn.Linear(12 * 4 * 4, 100), nn.Tanh(), nn.Linear(100, CLASSES))


ckpt_dir = tempfile.mkdtemp(prefix="zoo_lenet_ckpt_")
net = lenet()
opt = TFOptimizer.from_loss(nn.CrossEntropyLoss(), Adam(lr=1e-3), model=net, dataset=TFDataset.from_ndarrays((images[:split], labels[:split]), batch_size=256),
                            model_dir=ckpt_dir)
opt.optimize(end_trigger=MaxEpoch(int(os.environ.get("EPOCHS", 3))), checkpoint_trigger=None)
path = os.path.join(ckpt_dir, "lenet.pt")
torch.save(net.state_dict(), path)                                # train.py's `saver.save(sess, "/tmp/lenet/model")`

restored = lenet()
restored.load_state_dict(torch.load(path))                        # evaluate.py's `saver.restore(sess, "/tmp/lenet/model")`


class Correct(nn.Module):
    """The `correct = tf.nn.in_top_k(logits, labels, 1)` output of the reference's graph: 1.0 where the prediction is right."""

    def __init__(self, model):
        super().__init__()
        self.model = model

    def forward(self, x, y):
        return (self.model(x).argmax(-1) == y.reshape(-1)).float()


pairs = TFDataset.from_ndarrays(((images[split:], labels[split:]), None), batch_per_thread=20)      # two graph inputs, no label
hits = TFPredictor.from_outputs(Correct(restored), dataset=pairs).predict()
print("evaluated %d records in batches of 20: accuracy %.3f" % (len(hits), float(np.mean(hits))))
