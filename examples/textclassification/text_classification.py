"""Text classification (news20-shaped): TextSet pipeline -> TextClassifier with a CNN encoder."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.feature.text import LocalTextSet
from zoo.models.textclassification import TextClassifier

rng = np.random.RandomState(0)
topics = [["space", "orbit", "nasa", "launch", "moon"], ["hockey", "team", "game", "season", "goal"], ["windows", "driver", "file", "disk", "card"]]
filler = ["the", "a", "of", "and", "is", "in", "it", "for"]
texts, labels = [], []
for i in range(900):
    k = i % 3
    words = list(rng.choice(topics[k], 6)) + list(rng.choice(filler, 10))
    rng.shuffle(words)
    texts.append(" ".join(words))
    labels.append(k)
ts = LocalTextSet(texts, labels).tokenize().normalize().word2idx(remove_topN=0, max_words_num=100).shape_sequence(20).generate_sample()
x, y = ts.to_arrays()
model = TextClassifier(3, vocab_size=max(ts.get_word_index().values()) + 1, embed_dim=32, sequence_length=20, encoder="cnn", encoder_output_dim=64)
model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
model.fit(x[:700], y[:700], batch_size=100, nb_epoch=5, validation_data=(x[700:], y[700:]))
print("[loss, accuracy] on held-out texts:", model.evaluate(x[700:], y[700:], batch_size=100))
