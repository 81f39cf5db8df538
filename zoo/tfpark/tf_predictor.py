"""`zoo.tfpark.tf_predictor` (reference: pyzoo/zoo/tfpark/tf_predictor.py:27-90)."""
from .unsupported import TFPredictor  # noqa: F401
