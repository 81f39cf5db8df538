"""`zoo.tfpark.zoo_optimizer` (reference: pyzoo/zoo/tfpark/zoo_optimizer.py:22-140)."""
from .unsupported import ZooOptimizer  # noqa: F401


class FakeOptimMethod:
    """Placeholder OptimMethod the reference hands to BigDL when the real update happens inside the TF graph.  Here it is an
    optimizer description that changes nothing (learning rate 0)."""

    def __init__(self):
        self.learningrate = 0.0

    def to_opt_config(self):
        from ..pipeline.api.keras.optimizers import to_opt_config

        return to_opt_config({"name": "sgd", "lr": 0.0}) if False else None
