"""Base classes of the Keras-style API (reference: pyzoo/zoo/pipeline/api/keras/base.py:22-110).  In the reference these are
py4j handles (`ZooKerasCreator` builds the JVM object named after the Python class, `ZooCallable` makes it callable on
Variables, `InferShape` asks the JVM for shapes); here a layer IS a torch module that does all three."""
from .layers.base import ZooKerasLayer


class ZooKerasCreator:
    """Name compatibility: objects are created directly in Python, there is no JVM constructor to dispatch to."""

    def jvm_class_constructor(self):
        return "createZooKeras" + type(self).__name__


class ZooCallable:
    """Calling a layer on a `Variable` (or a list of them) returns a `Variable`: implemented by `ZooKerasLayer.__call__`."""


class InferShape:
    """`get_input_shape()` / `get_output_shape()` (batch dimension reported as None): implemented by `ZooKerasLayer`."""

    def get_input_shape(self):
        return ZooKerasLayer.get_input_shape(self)

    def get_output_shape(self):
        return ZooKerasLayer.get_output_shape(self)


__all__ = ["ZooKerasCreator", "ZooCallable", "InferShape", "ZooKerasLayer"]
