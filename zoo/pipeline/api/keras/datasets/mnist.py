import gzip
import os

import numpy as np

from ._synthetic import warn

TRAIN_MEAN, TRAIN_STD, TEST_MEAN, TEST_STD = 0.13066047740239506 * 255, 0.3081078 * 255, 0.13251460696903547 * 255, 0.31048024 * 255


def _read32(bytestream):
    return np.frombuffer(bytestream.read(4), dtype=np.dtype(np.uint32).newbyteorder(">"))[0]


def extract_images(f):
    """idx3-ubyte(.gz) stream -> uint8 [n, rows, cols, 1]."""
    with gzip.GzipFile(fileobj=f) as bs:
        if _read32(bs) != 2051:
            raise ValueError("Invalid magic number in MNIST image file")
        n, r, c = _read32(bs), _read32(bs), _read32(bs)
        return np.frombuffer(bs.read(int(r * c * n)), dtype=np.uint8).reshape(n, r, c, 1)


def extract_labels(f):
    with gzip.GzipFile(fileobj=f) as bs:
        if _read32(bs) != 2049:
            raise ValueError("Invalid magic number in MNIST label file")
        n = _read32(bs)
        return np.frombuffer(bs.read(int(n)), dtype=np.uint8)


def read_data_sets(train_dir, data_type="train"):
    img = "train-images-idx3-ubyte.gz" if data_type == "train" else "t10k-images-idx3-ubyte.gz"
    lab = "train-labels-idx1-ubyte.gz" if data_type == "train" else "t10k-labels-idx1-ubyte.gz"
    pi, pl = os.path.join(train_dir, img), os.path.join(train_dir, lab)
    if os.path.exists(pi) and os.path.exists(pl):
        with open(pi, "rb") as f:
            x = extract_images(f)
        with open(pl, "rb") as f:
            y = extract_labels(f)
        return x, y
    warn("mnist", train_dir)
    n = 60000 if data_type == "train" else 10000
    rng = np.random.RandomState(0 if data_type == "train" else 1)
    y = rng.randint(0, 10, n).astype(np.uint8)
    x = rng.randint(0, 64, (n, 28, 28, 1)).astype(np.uint8)
    for k in range(10):
        x[y == k, 2 * k + 4, :, 0] += 150
    return x, y


def load_data(location="/tmp/.zoo/dataset/mnist"):
    (x_train, y_train), (x_test, y_test) = read_data_sets(location, "train"), read_data_sets(location, "test")
    return (x_train, y_train), (x_test, y_test)
