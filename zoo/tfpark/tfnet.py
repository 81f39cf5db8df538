"""`zoo.tfpark.tfnet` (reference: pyzoo/zoo/tfpark/tfnet.py:35-302)."""
import numpy as np

from ..feature.common import Sample
from .unsupported import TFNet  # noqa: F401


def to_sample_rdd(x, y, sc=None, num_slices=None):
    """ndarrays -> a list ("RDD") of Samples, one per row."""
    xs = x if isinstance(x, (list, tuple)) else [x]
    ys = None if y is None else (y if isinstance(y, (list, tuple)) else [y])
    n = len(xs[0])
    out = []
    for i in range(n):
        f = [np.asarray(t[i]) for t in xs]
        l = None if ys is None else [np.asarray(t[i]) for t in ys]
        out.append(Sample(f[0] if len(f) == 1 else f, None if l is None else (l[0] if len(l) == 1 else l)))
    return out
