"""`VanillaLSTM`, the Keras-recipe forecaster (reference: pyzoo/zoo/automl/model/VanillaLSTM.py:30-200: LSTM(lstm_1_units=20,
return_sequences) -> Dropout(dropout_1=0.2) -> LSTM(lstm_2_units=10) -> Dropout(dropout_2=0.2) -> Dense(future_seq_len), RMSprop(lr),
batch_size 1024, Monte-Carlo dropout for the uncertainty estimate).  The architecture, defaults and optimizer are the reference's; the
two recurrent layers run on the persistent LSTM kernel (K6) and the training loop is `TorchForecastModel`'s."""
import torch

from .VanillaLSTM_pytorch import TwoLayerLSTM, VanillaLSTMPytorch


class VanillaLSTM(VanillaLSTMPytorch):
    def _build(self, config):
        return TwoLayerLSTM(self.feature_num, int(config.get("lstm_1_units", 20)), float(config.get("dropout_1", 0.2)),
                            int(config.get("lstm_2_units", 10)), float(config.get("dropout_2", 0.2)), self.future_seq_len)

    def _make_optimizer(self):
        # keras.optimizers.RMSprop defaults: rho 0.9, epsilon 1e-7
        return torch.optim.RMSprop(self.model.parameters(), lr=self.lr, alpha=0.9, eps=1e-7)

    def _get_optional_parameters(self):
        return {"lstm_1_units", "dropout_1", "lstm_2_units", "dropout_2", "lr", "batch_size", "epochs", "metric"}
