"""`zoo.pipeline.api.net.tf_net` (reference: pyzoo/zoo/pipeline/api/net/tf_net.py): TFNet runs a frozen TensorFlow graph."""
from .tfnet import TFNet  # noqa: F401
