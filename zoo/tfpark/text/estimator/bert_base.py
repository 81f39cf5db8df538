"""`bert_base` names (reference: pyzoo/zoo/tfpark/text/estimator/bert_base.py:21-150)."""
from . import BERTBaseEstimator, BertEncoder, bert_input_fn, load_bert_config  # noqa: F401


def bert_model(features, labels, mode, params):
    """Build the encoder for `params["bert_config_file"]` and run it on the feature dict; returns an object with
    `get_pooled_output()` / `get_sequence_output()` like Google's BertModel."""
    enc = BertEncoder(params.get("bert_config_file"))
    enc.train(mode == "train")
    seq, pooled = enc(features["input_ids"], features.get("input_mask"), features.get("token_type_ids"))

    class _Out:
        encoder = enc

        @staticmethod
        def get_pooled_output():
            return pooled

        @staticmethod
        def get_sequence_output():
            return seq

    return _Out
