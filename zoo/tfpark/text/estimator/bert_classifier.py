"""`bert_classifier` names (reference: pyzoo/zoo/tfpark/text/estimator/bert_classifier.py)."""
from . import BERTClassifier, _ClassifierNet  # noqa: F401


def make_bert_classifier_model_fn(optimizer):
    """The reference returns a TF model_fn closed over the optimizer; here: a factory `(params) -> (network, optimizer)`
    for the same head, usable with `BERTBaseEstimator`."""
    def model_fn(params):
        args = [params.get("bert_config_file")] + [params["num_classes"]]
        return _ClassifierNet(*args), optimizer

    return model_fn
