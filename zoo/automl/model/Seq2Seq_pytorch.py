"""LSTM encoder-decoder forecaster (reference: pyzoo/zoo/automl/model/Seq2Seq_pytorch.py:31-387; keras twin
Seq2Seq.py).  Encoder/Decoder keep the reference's sub-module names (`rnn`, `fc_out`); decoding is
autoregressive over `future_seq_len` steps starting from the last observed target value."""
from __future__ import annotations

import torch
import torch.nn as nn

from ...ops.rnn import run as rnn_run
from .torch_base import TorchForecastModel


def _init_rnn(rnn):
    for name, param in rnn.named_parameters():
        if "bias" in name:
            nn.init.constant_(param, 0.0)
        elif "weight_ih" in name:
            nn.init.xavier_normal_(param)
        elif "weight_hh" in name:
            nn.init.orthogonal_(param)


class Encoder(nn.Module):
    def __init__(self, input_dim, hidden_dim, layer_num, dropout):
        super().__init__()
        self.rnn = nn.LSTM(input_dim, hidden_dim, layer_num, dropout=dropout if layer_num > 1 else 0.0, batch_first=True)
        _init_rnn(self.rnn)

    def forward(self, input_seq):
        _, (hidden, cell) = rnn_run(self.rnn, input_seq.float())
        return hidden, cell


class Decoder(nn.Module):
    def __init__(self, output_dim, hidden_dim, layer_num, dropout):
        super().__init__()
        self.rnn = nn.LSTM(output_dim, hidden_dim, layer_num, dropout=dropout if layer_num > 1 else 0.0, batch_first=True)
        self.fc_out = nn.Linear(hidden_dim, output_dim)
        _init_rnn(self.rnn)

    def forward(self, decoder_input, hidden, cell):
        output, (hidden, cell) = rnn_run(self.rnn, decoder_input.unsqueeze(1), (hidden.contiguous(), cell.contiguous()))
        return self.fc_out(output[:, 0]), hidden, cell


class Seq2Seq(nn.Module):
    def __init__(self, encoder, decoder, target_seq_len=1):
        super().__init__()
        self.encoder, self.decoder, self.target_seq_len = encoder, decoder, target_seq_len

    def forward(self, source, target_seq=None, teacher_forcing_ratio=0.0):
        hidden, cell = self.encoder(source)
        decoder_input = source[:, -1, :1].float()            # last observed (scaled) target value
        outs = []
        for t in range(self.target_seq_len):
            out, hidden, cell = self.decoder(decoder_input, hidden, cell)
            outs.append(out)
            use_truth = target_seq is not None and self.training and torch.rand(1).item() < teacher_forcing_ratio
            decoder_input = target_seq[:, t:t + 1] if use_truth else out
        return torch.cat(outs, dim=1)


class Seq2SeqPytorch(TorchForecastModel):
    def _get_configs(self, config):
        super()._get_configs(config)
        self.latent_dim = int(config.get("latent_dim", 128))
        self.layer_num = int(config.get("layer_num", 2))
        self.dropout = float(config.get("dropout", 0.2))

    def _build(self, config):
        enc = Encoder(self.feature_num, self.latent_dim, self.layer_num, self.dropout)
        dec = Decoder(1, self.latent_dim, self.layer_num, self.dropout)
        return Seq2Seq(enc, dec, self.future_seq_len)

    def _get_optional_parameters(self):
        return {"latent_dim", "layer_num", "dropout", "lr", "batch_size"}


LSTMSeq2Seq = Seq2SeqPytorch
