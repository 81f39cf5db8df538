"""TFNet inference (reference example: pyzoo/zoo/examples/tensorflow/tfnet/predict.py: load a frozen TensorFlow detection graph
with `TFNet.from_export_folder`, wrap an ImageSet, `model.predict_image(...)`).  A frozen `GraphDef` is written here with
`export_tf` (the same artefacts the reference's `export_tf` produces: frozen_inference_graph.pb + graph_meta.json), then read back
by the framework's own protobuf reader and run as a torch module on the GPU -- no TensorFlow at either end."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.pipeline.api.net import TFNet
from zoo.util.tf import export_tf

torch.manual_seed(0)
cnn = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.BatchNorm2d(8), nn.ReLU(), nn.MaxPool2d(2), nn.Conv2d(8, 16, 3, stride=2), nn.ReLU(),
                    nn.Flatten(), nn.Linear(16 * 7 * 7, 10), nn.Softmax(-1)).eval()
folder = tempfile.mkdtemp()
export_tf(cnn, folder, torch.zeros(1, 3, 32, 32))
print("exported:", sorted(os.listdir(folder)))

net = TFNet.from_export_folder(folder)
images = torch.rand(6, 3, 32, 32)
nhwc = images.permute(0, 2, 3, 1).contiguous()             # TensorFlow graphs take NHWC
dev = "cuda" if torch.cuda.is_available() and not os.environ.get("ZOO_FORCE_CPU") else "cpu"
with torch.no_grad():
    got = net.to(dev)(nhwc.to(dev)).cpu()
    want = cnn(images)
print("predicted %d images, classes %s, max |TFNet - torch| = %.2e" % (len(got), got.argmax(1).tolist(), float((got - want).abs().max())))
