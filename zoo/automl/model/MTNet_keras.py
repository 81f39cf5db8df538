"""`MTNetKeras` / `AttentionRNNWrapper` names of the reference's tf.keras MTNet (pyzoo/zoo/automl/model/MTNet_keras.py:54-600).
MTNetKeras is the torch MTNet here; AttentionRNNWrapper (Luong-style attention with input feeding around an RNN layer) is
provided as a torch module over a GRU/LSTM cell."""
import torch
import torch.nn as nn

from .MTNet import MTNet


class MTNetKeras(MTNet):
    pass


class AttentionRNNWrapper(nn.Module):
    """After every step an attention vector over the WHOLE input series is computed from the current RNN output, the attended
    input is concatenated to the next step's input, and a linear map brings it back to the input space before the cell."""

    def __init__(self, layer, weight_initializer="glorot_uniform"):
        super().__init__()
        if not isinstance(layer, (nn.GRU, nn.LSTM, nn.RNN)):
            raise AssertionError("layer must be a torch RNN layer (nn.GRU / nn.LSTM / nn.RNN)")
        if not layer.batch_first or layer.num_layers != 1 or layer.bidirectional:
            raise ValueError("AttentionRNNWrapper wraps a single-layer, unidirectional, batch_first RNN")
        self.kind = type(layer).__name__
        cell = {"GRU": nn.GRUCell, "LSTM": nn.LSTMCell, "RNN": nn.RNNCell}[self.kind]
        self.cell = cell(layer.input_size, layer.hidden_size)
        d, h = layer.input_size, layer.hidden_size
        self.w1, self.w2 = nn.Linear(d, h, bias=False), nn.Linear(h, h, bias=False)      # attention energies
        self.v = nn.Linear(h, 1, bias=False)
        self.w3 = nn.Linear(2 * d, d)                                                    # input feeding
        for m in (self.w1, self.w2, self.v, self.w3):
            nn.init.xavier_uniform_(m.weight)
        self.return_sequences = False

    def forward(self, x):
        B, T, D = x.shape
        h = x.new_zeros(B, self.cell.hidden_size)
        c = x.new_zeros(B, self.cell.hidden_size)
        keys = self.w1(x)
        attended = x.new_zeros(B, D)
        outs = []
        for t in range(T):
            inp = self.w3(torch.cat([x[:, t], attended], -1))
            if self.kind == "LSTM":
                h, c = self.cell(inp, (h, c))
            else:
                h = self.cell(inp, h)
            e = self.v(torch.tanh(keys + self.w2(h)[:, None, :])).squeeze(-1)
            attended = (torch.softmax(e, -1)[:, :, None] * x).sum(1)
            outs.append(h)
        return torch.stack(outs, 1) if self.return_sequences else h
