"""Cluster Serving quick start: start an engine in this process, enqueue images, query the results."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch.nn as nn

from zoo.serving import ClusterServing, InputQueue, OutputQueue

model = nn.Sequential(nn.Conv2d(3, 8, 3, stride=4), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(8, 10), nn.Softmax(-1))
conf = {"data": {"image_shape": "3,224,224", "filter": "topN(3)"}, "params": {"batch_size": 16}}
serving = ClusterServing(conf, model=model).start()
inq, outq = InputQueue(conf, broker=serving.db), OutputQueue(conf, broker=serving.db)
rng = np.random.RandomState(0)
for i in range(32):
    inq.enqueue_image(f"img-{i}", rng.randint(0, 256, (300, 400, 3), dtype=np.uint8))
serving.wait_until(32)
print("img-7 ->", outq.query("img-7"))
print(len(outq.dequeue()), "results dequeued;", serving.stats())
serving.close()
