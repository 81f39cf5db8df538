from .estimator import Estimator  # noqa: F401
from .local_estimator import LocalEstimator  # noqa: F401
