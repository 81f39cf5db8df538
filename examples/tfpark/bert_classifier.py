"""BERTClassifier (reference example: pyzoo/zoo/examples/tensorflow/tfpark/estimator/... bert_classifier usage in docs:
`bert_input_fn(rdd, max_seq_length, batch_size)` -> `estimator.train(input_fn, steps)` -> `evaluate` -> `predict`)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch

from zoo.tfpark.text import BERTClassifier, bert_input_fn

L, n = 16, 512
rng = np.random.RandomState(0)
records = []
for _ in range(n):
    ids = rng.randint(5, 100, L)
    ids[0] = 1                                                            # [CLS]
    label = int((ids[1:] % 2 == 0).sum() > (L - 1) / 2)
    records.append(({"input_ids": ids, "input_mask": np.ones(L, np.int64), "token_type_ids": np.zeros(L, np.int64)}, label))
config = dict(vocab_size=100, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, intermediate_size=128, max_position_embeddings=32)
est = BERTClassifier(2, config, optimizer=lambda p: torch.optim.AdamW(p, lr=1e-3), device="cpu")
train_fn = bert_input_fn(records[:448], L, 32)
test_fn = bert_input_fn(records[448:], L, 32)
est.train(train_fn, steps=int(os.environ.get("STEPS", 200)))
print("eval:", est.evaluate(test_fn, ["acc", "loss"]))
probs = est.predict(bert_input_fn([r[0] for r in records[448:452]], L, 4))
print("probabilities of 4 records:", np.round(np.stack(probs), 3).tolist())
