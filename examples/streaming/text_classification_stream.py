"""Streaming inference: records arrive on a stream, a model server classifies micro-batches, results are read back
(reference examples: pyzoo/zoo/examples/streaming/textclassification -- Spark Streaming socketTextStream -> TextClassifier;
here the stream is the in-process broker that Cluster Serving uses, consumed with continuous batching)."""
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch

from zoo.models.textclassification import TextClassifier
from zoo.serving import ClusterServing, InputQueue, OutputQueue

torch.manual_seed(0)
seq_len, vocab = 20, 50
model = TextClassifier(class_num=3, sequence_length=seq_len, encoder="cnn", encoder_output_dim=32, vocab_size=vocab, embed_dim=16)
conf = {"data": {"filter": "topN(1)"}, "params": {"batch_size": 32}}
srv = ClusterServing(conf, model=model, device="cpu").start()
inq, outq = InputQueue(conf, broker=srv.db), OutputQueue(conf, broker=srv.db)
rng = np.random.RandomState(0)
n = 200
for i in range(n):                                   # the "socket": one record at a time
    inq.enqueue_tensor(f"line-{i}", rng.randint(1, vocab, seq_len).astype(np.int64))
assert srv.wait_until(n, 60)
time.sleep(0.05)
res = outq.dequeue()
print(len(res), "lines classified; first three:", [(k, res[k]) for k in sorted(res)[:3]])
srv.close()
