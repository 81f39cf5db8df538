"""Shape helpers of the Keras-style API (reference: pyzoo/zoo/pipeline/api/utils.py:24-40)."""


def toMultiShape(shape):
    if any(isinstance(i, list) for i in shape):
        return shape
    if any(isinstance(i, tuple) for i in shape):
        return [list(s) for s in shape]
    if isinstance(shape, tuple):
        return [list(shape)]
    return [shape]


def remove_batch(shape):
    if any(isinstance(i, (list, tuple)) for i in shape):
        return [remove_batch(s) for s in shape]
    return list(shape[1:])
