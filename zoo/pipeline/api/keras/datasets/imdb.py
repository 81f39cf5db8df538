import os
import pickle

import numpy as np

from ._synthetic import shuffle_by_seed, warn  # noqa: F401


def download_imdb(dest_dir):
    raise OSError("no network on this box: place imdb_full.pkl under " + dest_dir)


def _synthetic(n, vocab, seed):
    rng = np.random.RandomState(seed)
    y = rng.randint(0, 2, n)
    x = [list(rng.randint(3, vocab // 2, rng.randint(20, 200)) + (vocab // 2) * (yy if rng.rand() < 0.6 else 0)) for yy in y]
    return x, list(y)


def load_data(dest_dir="/tmp/.zoo/dataset", nb_words=None, oov_char=2):
    path = os.path.join(dest_dir, "imdb_full.pkl")
    if os.path.exists(path):
        with open(path, "rb") as f:
            (x_train, y_train), (x_test, y_test) = pickle.load(f)
    else:
        warn("imdb", dest_dir)
        vocab = nb_words or 5000
        (x_train, y_train), (x_test, y_test) = _synthetic(25000, vocab, 0), _synthetic(25000, vocab, 1)
    shuffle_by_seed([x_train, y_train, x_test, y_test])
    if nb_words:
        x_train = [[oov_char if w >= nb_words else w for w in x] for x in x_train]
        x_test = [[oov_char if w >= nb_words else w for w in x] for x in x_test]
    return (x_train, y_train), (x_test, y_test)


def get_word_index(dest_dir="/tmp/.zoo/dataset", ):
    path = os.path.join(dest_dir, "imdb_word_index.pkl")
    if os.path.exists(path):
        with open(path, "rb") as f:
            return pickle.load(f)
    return {f"w{i}": i for i in range(3, 5000)}
