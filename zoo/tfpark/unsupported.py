"""Old home of the TFPark names that used to be placeholders (reference: tf_optimizer.py:342-709, estimator.py:30-332,
tfnet.py:52-302, tf_predictor.py, zoo_optimizer.py:27).  All of them are implemented in their own modules now; this module only
forwards the names for code that imported them from here."""
from __future__ import annotations

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


from .zoo_optimizer import ZooOptimizer  # noqa: E402,F401
