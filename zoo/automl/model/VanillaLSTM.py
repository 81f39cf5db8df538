"""`VanillaLSTM` of the reference is a tf.keras model (pyzoo/zoo/automl/model/VanillaLSTM.py:30-200: two LSTM layers with
dropout and a Dense head, config keys lstm_1_units / dropout_1 / lstm_2_units / dropout_2 / lr / batch_size / epochs).
Here the same name is the torch model with the persistent-kernel LSTM (K6)."""
from .VanillaLSTM_pytorch import VanillaLSTMPytorch


class VanillaLSTM(VanillaLSTMPytorch):
    pass
