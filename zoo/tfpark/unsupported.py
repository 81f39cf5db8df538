"""TFPark pieces that are TensorFlow-graph machinery (reference: tf_optimizer.py:342-709, estimator.py:30-332,
tfnet.py:52-302, tf_predictor.py, gan/gan_estimator.py:28, zoo_optimizer.py:27).  No TF runtime exists here."""
from __future__ import annotations

_MSG = ("{name} drives a TensorFlow graph and needs the TensorFlow runtime, which is not part of zoo-b200 (SURVEY 7.3). "
        "Use zoo.pipeline.estimator.Estimator.from_torch / zoo.tfpark.KerasModel over a torch model instead.")


def _stub(name, methods=()):
    def __init__(self, *a, **kw):
        raise NotImplementedError(_MSG.format(name=name))

    ns = {"__init__": __init__, "__doc__": _MSG.format(name=name), "__zoo_stub__": True}
    for m in methods:
        ns[m] = staticmethod(lambda *a, _n=f"{name}.{m}", **kw: (_ for _ in ()).throw(NotImplementedError(_MSG.format(name=_n))))
    return type(name, (), ns)


from ..pipeline.api.net.tfnet import TFNet  # noqa: E402,F401  (real: frozen graphs are imported into torch, see zoo.tfpark.graphdef)


def __getattr__(name):      # TFOptimizer / TFEstimator / TFPredictor are real now and live in their own modules
    if name == "TFOptimizer":
        from .tf_optimizer import TFOptimizer
        return TFOptimizer
    if name == "TFEstimator":
        from .estimator import TFEstimator
        return TFEstimator
    if name == "TFPredictor":
        from .tf_predictor import TFPredictor
        return TFPredictor
    raise AttributeError(name)


class ZooOptimizer:
    """In the reference a wrapper that makes a tf.train.Optimizer usable by TFOptimizer.  Here: identity over a
    zoo / torch optimizer description, so code that wraps its optimizer keeps working."""

    def __init__(self, optimizer, name=None):
        self.optimizer = optimizer

    def __getattr__(self, item):
        return getattr(self.optimizer, item)
