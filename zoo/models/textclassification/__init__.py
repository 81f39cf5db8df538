from .text_classifier import TextClassifier  # noqa: F401
