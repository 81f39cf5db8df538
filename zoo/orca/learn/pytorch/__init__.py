from .estimator import Estimator  # noqa: F401
