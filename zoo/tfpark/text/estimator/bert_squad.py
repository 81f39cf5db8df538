"""`bert_squad` names (reference: pyzoo/zoo/tfpark/text/estimator/bert_squad.py)."""
from . import BERTSQuAD, _SquadNet  # noqa: F401


def make_bert_squad_model_fn(optimizer):
    """The reference returns a TF model_fn closed over the optimizer; here: a factory `(params) -> (network, optimizer)`
    for the same head, usable with `BERTBaseEstimator`."""
    def model_fn(params):
        args = [params.get("bert_config_file")] + []
        return _SquadNet(*args), optimizer

    return model_fn
