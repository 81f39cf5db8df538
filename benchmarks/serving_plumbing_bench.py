"""Throughput of the Cluster Serving plumbing (bulk producer -> packed batch read -> model -> native top-N result writer) with a trivial
model, producer and engine measured separately.  Runs on the CPU: it measures the broker / engine code, not the GPU.
usage: python benchmarks/serving_plumbing_bench.py"""
import sys, time
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.nn as nn
from zoo.serving import ClusterServing, InputQueue, OutputQueue
from zoo.native.rt import Broker

class Score(nn.Module):
    def forward(self, x):
        x = x.float()
        return torch.softmax(torch.stack([x[:, 0], x[:, 1], x[:, 0] - x[:, 1]], 1) * 1e-6, -1)

conf = {"data": {"filter": "topN(1)"}, "params": {"batch_size": 4096}}
total = 400000
db = Broker()
q = InputQueue(conf, broker=db)
rng = np.random.RandomState(0)
q.enqueue_tensor("r0", np.array([1, 1], dtype=np.int64))
t0 = time.time()
for s0 in range(1, total, 4096):
    n = min(4096, total - s0)
    ids = np.stack([rng.randint(1, 1_000_001, n), rng.randint(1, 100_001, n)], 1).astype(np.int64)
    q.enqueue_tensors([f"r{s0 + k}" for k in range(n)], ids)
print(f"enqueue only: {total / (time.time() - t0):,.0f} rec/s")
srv = ClusterServing(conf, model=Score(), device="cpu", broker=db) if "broker" in ClusterServing.__init__.__code__.co_varnames else None
if srv is None:
    print("no broker arg"); sys.exit()
t0 = time.time(); srv.start()
assert srv.wait_until(total, 300)
print(f"engine only: {total / (time.time() - t0):,.0f} rec/s")
srv.close()
