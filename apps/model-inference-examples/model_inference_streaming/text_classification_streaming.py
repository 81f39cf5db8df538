"""Streaming model inference (reference: apps/model-inference-examples/model-inference-flink: TextClassificationStreaming.scala
reads lines from a socket source, maps them through a `TextClassificationInferenceModel extends InferenceModel` inside a Flink
`RichMapFunction` (model loaded in `open`, released in `close`) and prints the predictions).

The dataflow engine here is the framework's own stream broker (the one Cluster Serving runs on): a source thread appends lines to a
stream, a map operator with `open / map / close` pulls micro-batches through a consumer group, and a sink prints.  The operator
batches whatever has arrived (continuous batching) instead of Flink's record-at-a-time map."""
import os
import sys
import tempfile
import threading
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "text_classification_inference"))
sys.path.insert(0, os.path.join(HERE, "..", "..", ".."))
import numpy as np
import torch
import torch.nn as nn
from text_classification_model import TextClassificationModel

from zoo.native.rt import Broker


class ModelPredictionMapFunction:
    """RichMapFunction: `open` loads the model once per operator instance, `map` scores a micro-batch, `close` releases it."""

    def __init__(self, emb_path, model_path, sequence_length):
        self.args = (emb_path, model_path, sequence_length)

    def open(self):
        emb, path, seq = self.args
        self.model = TextClassificationModel(1, 10, seq, emb)
        self.model.loadPyTorch(path)

    def map(self, lines):
        results = self.model.predict([[self.model.preprocess(t)] for t in lines])
        return [(t, int(r[0].to_ndarray().argmax()), float(r[0].to_ndarray().max())) for t, r in zip(lines, results)]

    def close(self):
        self.model.release()


if __name__ == "__main__":
    work = tempfile.mkdtemp(prefix="zoo_stream_textcls_")
    rng = np.random.RandomState(0)
    vocab = ["the", "a", "of", "to", "and", "in", "is", "it", "that", "for"] + ["space%d" % k for k in range(8)] + ["auto%d" % k for k in range(8)]
    emb = os.path.join(work, "glove.txt")
    with open(emb, "w") as f:
        for w in vocab:
            f.write(w + " " + " ".join("%.4f" % v for v in rng.randn(16)) + "\n")
    tp = TextClassificationModel(1, 10, 12, emb)
    make = lambda k: " ".join(rng.permutation(list(rng.choice(vocab[10 + 8 * k:18 + 8 * k], 5)) + list(rng.choice(vocab[:10], 4))))
    texts, labels = zip(*[(make(k % 2), k % 2) for k in range(300)])
    x, y = torch.from_numpy(np.stack([tp.preprocess(t).to_ndarray() for t in texts])), torch.tensor(labels)
    torch.manual_seed(0)
    net = nn.Sequential(nn.Flatten(), nn.Linear(12 * 16, 32), nn.ReLU(), nn.Linear(32, 2), nn.Softmax(-1))
    opt = torch.optim.Adam(net.parameters(), 1e-2)
    for _ in range(int(os.environ.get("TRAIN_STEPS", 40))):
        opt.zero_grad()
        nn.functional.nll_loss(torch.log(net(x) + 1e-9), y).backward()
        opt.step()
    model_path = os.path.join(work, "model.pt")
    torch.save(net.eval(), model_path)

    broker = Broker()
    total = 40
    truth = {}

    def source():                                              # the socket text stream
        for k in range(total):
            line = make(k % 2)
            truth[line] = k % 2
            broker.xadd("lines", {"text": line})
            time.sleep(0.002)

    op = ModelPredictionMapFunction(emb, model_path, 12)
    op.open()
    broker.xgroup_create("lines", "inference", id="0", mkstream=True)
    threading.Thread(target=source, daemon=True).start()
    seen = correct = batches = 0
    deadline = time.time() + 30
    while seen < total and time.time() < deadline:
        got = broker.xreadgroup("inference", "op-0", {"lines": ">"}, count=16, block=100)
        entries = [fields for _stream, items in (got or []) for _id, fields in items]
        if not entries:
            continue
        batches += 1
        as_text = lambda v: v.decode() if isinstance(v, (bytes, bytearray)) else v
        for text, cls_, prob in op.map([as_text(e.get("text", e.get(b"text"))) for e in entries]):         # the sink
            seen += 1
            correct += int(cls_ == truth[text])
            if seen <= 3:
                print("%-70s -> class %d (%.2f)" % (text[:70], cls_, prob))
    op.close()
    print("%d lines classified in %d micro-batches, %d correct" % (seen, batches, correct))
