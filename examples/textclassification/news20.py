"""Data helpers of the text-classification example (reference: pyzoo/zoo/examples/textclassification/news20.py: `get_news20(dir)` ->
list of (text, label) from the 20 Newsgroups folders, `get_glove_w2v(dir, dim)` -> word -> vector from the GloVe text file; both
download their archives on first use).  Offline here: when the folder is missing the synthetic corpora of
`scripts/data/gen_synthetic.py` are generated in the same layouts, so `text_classification.py --data_path DIR` runs either way."""
import os
import subprocess
import sys

import numpy as np

_GEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "scripts", "data", "gen_synthetic.py")
CLASS_NUM = 20


def _ensure(kind, path):
    if not os.path.isdir(path) or not os.listdir(path):
        subprocess.run([sys.executable, _GEN, kind, path], check=True, stdout=subprocess.DEVNULL)
    return path


def get_news20(source_dir="/tmp/news20/"):
    """[(text, label)], labels 0..n-1 in the sorted order of the category folders (the 20news-18828 layout: one folder per group,
    one file per post)."""
    root = _ensure("news20", os.path.join(source_dir, "20news-18828") if os.path.isdir(os.path.join(source_dir, "20news-18828")) else source_dir)
    texts = []
    for label, name in enumerate(sorted(d for d in os.listdir(root) if os.path.isdir(os.path.join(root, d)))):
        folder = os.path.join(root, name)
        for fname in sorted(os.listdir(folder)):
            with open(os.path.join(folder, fname), encoding="latin-1") as f:
                texts.append((f.read(), label))
    print("Found %s texts." % len(texts))
    return texts


def get_glove_w2v(source_dir="/tmp/news20/glove.6B/", dim=100):
    """word -> vector (list of floats) from glove.6B.<dim>d.txt."""
    folder = _ensure("glove", source_dir)
    path = os.path.join(folder, "glove.6B.%sd.txt" % dim)
    if not os.path.exists(path):                                  # the synthetic file is 50-d
        path = next(os.path.join(folder, f) for f in sorted(os.listdir(folder)) if f.startswith("glove.6B."))
    w2v = {}
    with open(path, encoding="utf-8") as f:
        for line in f:
            parts = line.rstrip().split(" ")
            w2v[parts[0]] = [float(v) for v in parts[1:]]
    return w2v


if __name__ == "__main__":
    import tempfile

    d = tempfile.mkdtemp()
    texts = get_news20(os.path.join(d, "news20"))
    w2v = get_glove_w2v(os.path.join(d, "glove.6B"), 50)
    print("classes: %d, vocabulary of the embedding file: %d, vector width: %d" % (len({l for _, l in texts}), len(w2v), len(next(iter(w2v.values())))))
