from .nn_classifier import NNClassifier, NNClassifierModel, NNEstimator, NNModel  # noqa: F401
from .nn_image_reader import NNImageReader  # noqa: F401
