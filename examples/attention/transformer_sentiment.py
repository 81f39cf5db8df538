"""Sentiment classification with a TransformerLayer encoder (reference example: pyzoo/zoo/examples/attention/transformer.py:
IMDB, TransformerLayer -> SelectTable(pooled) -> Dropout -> Dense(2, softmax)).  Synthetic "reviews": the label is whether
positive tokens outnumber negative ones."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.pipeline.api.keras.engine.topology import Model
from zoo.pipeline.api.keras.layers import Dense, Dropout, Input, SelectTable, TransformerLayer

max_features, max_len, n = 200, 24, 3000
rng = np.random.RandomState(0)
tokens = rng.randint(10, max_features, (n, max_len))
score = ((tokens % 7 == 0).sum(1) - (tokens % 7 == 1).sum(1))
y = (score > 0).astype(np.int64)
pos = np.tile(np.arange(max_len), (n, 1)) + max_features            # positions are extra vocabulary entries

tok_in, pos_in = Input(shape=(max_len,)), Input(shape=(max_len,))
enc = TransformerLayer.init(vocab=max_features + max_len, seq_len=max_len, n_block=2, n_head=4, hidden_size=64, hidden_drop=0.1, attn_drop=0.1,
                            embedding_drop=0.1, bidirectional=True)
pooled = SelectTable(1)(enc([tok_in, pos_in]))
out = Dense(2, activation="softmax")(Dropout(0.1)(pooled))
model = Model([tok_in, pos_in], out)
model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
split = 2500
model.fit([tokens[:split], pos[:split]], y[:split], batch_size=128, nb_epoch=int(os.environ.get("EPOCHS", 6)))
res = model.evaluate([tokens[split:], pos[split:]], y[split:], batch_size=250)
print("test loss %.4f  accuracy %.3f" % (res[0], res[1]))
