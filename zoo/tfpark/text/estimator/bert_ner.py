"""`bert_ner` names (reference: pyzoo/zoo/tfpark/text/estimator/bert_ner.py)."""
from . import BERTNER, _NerNet  # noqa: F401


def make_bert_ner_model_fn(optimizer):
    """The reference returns a TF model_fn closed over the optimizer; here: a factory `(params) -> (network, optimizer)`
    for the same head, usable with `BERTBaseEstimator`."""
    def model_fn(params):
        args = [params.get("bert_config_file")] + [params["num_entities"]]
        return _NerNet(*args), optimizer

    return model_fn
