"""Sentiment analysis (reference app: apps/sentiment-analysis/sentiment.ipynb -- IMDB reviews, GloVe embedding, CNN / LSTM / GRU
/ Bi-LSTM / CNN-LSTM heads built with the Keras-style API).  The TextSet pipeline is the reference's: tokenize -> normalize ->
word2idx -> shape_sequence -> generate_sample; reviews are generated from two word pools."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.feature.text import LocalTextSet
from zoo.pipeline.api.keras.layers import LSTM, Convolution1D, Dense, Dropout, Embedding, GlobalMaxPooling1D
from zoo.pipeline.api.keras.models import Sequential

rng = np.random.RandomState(0)
good, bad, filler = ["great", "wonderful", "loved", "brilliant", "superb"], ["awful", "boring", "hated", "terrible", "dull"], \
    ["the", "movie", "was", "plot", "acting", "and", "a", "really", "film", "it"]
texts, labels = [], []
for i in range(1200):
    lab = i % 2
    pool = good if lab else bad
    words = [rng.choice(pool) if rng.rand() < 0.25 else rng.choice(filler) for _ in range(rng.randint(15, 40))]
    texts.append(" ".join(words) + ("!" if lab else "."))
    labels.append(lab)
seq_len = 40
ts = LocalTextSet(texts, labels).tokenize().normalize().word2idx(remove_topN=0, max_words_num=5000).shape_sequence(seq_len).generate_sample()
x, _ = ts.to_arrays()
y = np.asarray(labels, dtype=np.int64)
vocab = int(x.max()) + 1
kind = os.environ.get("ENCODER", "cnn-lstm")
model = Sequential()
model.add(Embedding(vocab, 32, input_length=seq_len))
model.add(Dropout(0.2))
if kind in ("cnn", "cnn-lstm"):
    model.add(Convolution1D(32, 5, activation="relu"))
if kind == "cnn":
    model.add(GlobalMaxPooling1D())
else:
    model.add(LSTM(32))
model.add(Dense(2, activation="softmax"))
model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
model.fit(x[:1000], y[:1000], batch_size=64, nb_epoch=int(os.environ.get("EPOCHS", 4)))
res = model.evaluate(x[1000:], y[1000:], batch_size=100)
print("encoder %s: held-out loss %.4f accuracy %.3f" % (kind, res[0], res[1]))
