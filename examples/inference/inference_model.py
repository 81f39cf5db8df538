"""InferenceModel: several threads share one model through a pool of execution slots."""
import os
import sys
import threading

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch.nn as nn

from zoo.pipeline.inference import InferenceModel

model = InferenceModel(supported_concurrent_num=4).load_torch(nn.Sequential(nn.Linear(32, 64), nn.ReLU(), nn.Linear(64, 5), nn.Softmax(-1)))
x = np.random.rand(256, 32).astype(np.float32)
outs = {}
threads = [threading.Thread(target=lambda i=i: outs.__setitem__(i, model.predict(x))) for i in range(8)]
[t.start() for t in threads]
[t.join() for t in threads]
print(model, "records:", model.records, "all equal:", all(np.allclose(outs[0], o) for o in outs.values()))
