import os
import pickle

import numpy as np

from ._synthetic import shuffle_by_seed, warn  # noqa: F401


def download_reuters(dest_dir):
    raise OSError("no network on this box: place reuters.pkl under " + dest_dir)


def load_data(dest_dir="/tmp/.zoo/dataset", nb_words=None, oov_char=2, test_split=0.2):
    path = os.path.join(dest_dir, "reuters.pkl")
    if os.path.exists(path):
        with open(path, "rb") as f:
            x, y = pickle.load(f)
    else:
        warn("reuters", dest_dir)
        rng = np.random.RandomState(0)
        y = list(rng.randint(0, 46, 11228))
        x = [list(rng.randint(3, 2000, rng.randint(10, 150)) + 40 * yy) for yy in y]
    shuffle_by_seed([x, y])
    if nb_words:
        x = [[oov_char if w >= nb_words else w for w in s] for s in x]
    k = int(len(x) * (1 - test_split))
    return (x[:k], y[:k]), (x[k:], y[k:])


def get_word_index(dest_dir="/tmp/.zoo/dataset", ):
    path = os.path.join(dest_dir, "reuters_word_index.pkl")
    if os.path.exists(path):
        with open(path, "rb") as f:
            return pickle.load(f)
    return {f"w{i}": i for i in range(3, 4000)}
