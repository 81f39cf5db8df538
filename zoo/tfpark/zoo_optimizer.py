"""`zoo.tfpark.zoo_optimizer` (reference: pyzoo/zoo/tfpark/zoo_optimizer.py:22-140).

`ZooOptimizer` in the reference subclasses `tf.train.Optimizer` so that TFOptimizer can find the gradients an optimizer computes.
With no TensorFlow graph here it is a transparent wrapper: `TFOptimizer.from_loss(loss, ZooOptimizer(Adam(...)))` keeps working and
the wrapped description reaches the engine unchanged.  `FakeOptimMethod` is the optimizer the reference hands to the JVM when the
real update runs inside the TF graph: a description that leaves the weights where they are (SGD with learning rate 0)."""
from ..pipeline.api.keras.optimizers import SGD, OptimMethod


class FakeOptimMethod(SGD):
    def __init__(self):
        super().__init__(learningrate=0.0)


class ZooOptimizer(OptimMethod):
    def __init__(self, optimizer, name=None):
        if isinstance(optimizer, ZooOptimizer):
            optimizer = optimizer.optimizer
        self.optimizer, self.name = optimizer, name or type(optimizer).__name__

    def to_config(self):
        from ..pipeline.api.keras.optimizers import to_opt_config

        return to_opt_config(self.optimizer)

    def __getattr__(self, item):
        if item == "optimizer":
            raise AttributeError(item)
        return getattr(self.optimizer, item)
