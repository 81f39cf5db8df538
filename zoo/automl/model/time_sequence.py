"""TimeSequenceModel: picks the concrete model from the config (reference: pyzoo/zoo/automl/model/time_sequence.py:28-136):
`model` = "LSTM" (default when future_seq_len == 1) | "Seq2seq" (default otherwise) | "MTNet" | "TCN"."""
from __future__ import annotations

from .abstract import BaseModel
from .MTNet import MTNet
from .Seq2Seq_pytorch import Seq2SeqPytorch
from .TCN import TCNPytorch
from .Seq2Seq import LSTMSeq2Seq
from .VanillaLSTM import VanillaLSTM
from .VanillaLSTM_pytorch import VanillaLSTMPytorch

# "<name>_keras": the reference's Keras recipes (two stacked LSTMs / single-layer teacher-forced encoder-decoder, RMSprop)
MODELS = {"lstm_keras": VanillaLSTM, "seq2seq_keras": LSTMSeq2Seq, "lstm": VanillaLSTMPytorch, "vanillalstm": VanillaLSTMPytorch,
        "seq2seq": Seq2SeqPytorch, "lstmseq2seq": Seq2SeqPytorch,
          "mtnet": MTNet, "tcn": TCNPytorch}


class TimeSequenceModel(BaseModel):
    def __init__(self, check_optional_config=False, future_seq_len=None):
        self.check_optional_config, self.future_seq_len = check_optional_config, future_seq_len
        self.model, self.selected_model = None, None

    def _sel_model(self, config, verbose=0):
        name = config.get("model")
        if name is None:
            name = "LSTM" if (self.future_seq_len or 1) == 1 else "Seq2seq"
        key = str(name).lower()
        if key not in MODELS:
            raise ValueError(f"model {name!r} is not supported: choose from LSTM, Seq2seq, MTNet, TCN")
        self.selected_model = name
        self.model = MODELS[key](check_optional_config=self.check_optional_config, future_seq_len=self.future_seq_len or 1)
        return self.model

    def fit_eval(self, x, y, validation_data=None, mc=False, verbose=0, **config):
        if self.model is None:
            self._sel_model(config, verbose)
        return self.model.fit_eval(x, y, validation_data=validation_data, mc=mc, verbose=verbose, **config)

    def evaluate(self, x, y, metric=None):
        return self.model.evaluate(x, y, metric)

    def predict(self, x, mc=False):
        return self.model.predict(x, mc=mc)

    def predict_with_uncertainty(self, x, n_iter=100):
        return self.model.predict_with_uncertainty(x, n_iter)

    def save(self, model_path, config_path):
        self.model.save(model_path, config_path)
        from ..common.util import save_config

        save_config(config_path, {"model": self.selected_model})

    def restore(self, model_path, **config):
        self.future_seq_len = config.get("future_seq_len", self.future_seq_len)
        self._sel_model(config)
        self.model.restore(model_path, **config)
        return self
