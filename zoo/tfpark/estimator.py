"""`zoo.tfpark.estimator` (reference: pyzoo/zoo/tfpark/estimator.py:30-332).  `TFEstimator.from_model_fn` builds TF graphs;
the torch estimators with the same train / evaluate / predict surface are `zoo.tfpark.text.BERTBaseEstimator` and
`zoo.pipeline.estimator.Estimator.from_torch`."""
from .unsupported import TFEstimator  # noqa: F401
