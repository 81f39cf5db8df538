"""Seq2seq (reference: pyzoo/zoo/models/seq2seq/seq2seq.py:30-300, zoo/.../models/seq2seq/*.scala):
`RNNEncoder` / `RNNDecoder` (stacked lstm | gru | simplernn, optional embedding), `Bridge` (dense | densenonlinear
| custom layer, maps encoder states to decoder states), `Seq2seq(encoder, decoder, input_shape, output_shape,
bridge, generator)` with `fit` on (encoder input, decoder input) -> target and a greedy `infer` loop.
Recurrent layers run on kernel K6 (`zoo.ops.rnn`)."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from ...ops import rnn as _rnn
from ..common.zoo_model import ZooModel


def createRNN(rnn_type, nlayers, hidden_size, input_size=None):
    t = rnn_type.lower()
    cls = {"lstm": nn.LSTM, "gru": nn.GRU, "simplernn": nn.RNN}.get(t)
    if cls is None:
        raise ValueError("Only support lstm|gru|simplernn")
    return (t, int(nlayers), int(hidden_size), cls)


class _RNNStack(nn.Module):
    def __init__(self, rnns, embedding=None, input_shape=None):
        super().__init__()
        self.rnn_type, self.nlayers, self.hidden_size, cls = rnns
        self.embedding = embedding
        self.input_shape = input_shape
        in_dim = embedding.embedding_dim if isinstance(embedding, nn.Embedding) else (input_shape[-1] if input_shape else self.hidden_size)
        if embedding is not None and not isinstance(embedding, nn.Embedding):
            in_dim = getattr(embedding, "output_dim", in_dim)
        self.rnn = cls(in_dim, self.hidden_size, self.nlayers, batch_first=True)

    def _embed(self, x):
        return self.embedding(x.long()) if self.embedding is not None else x.float()

    def run(self, x, state=None):
        if isinstance(self.rnn, nn.RNN):
            return self.rnn(x, state)
        return _rnn.run(self.rnn, x, state)


class RNNEncoder(_RNNStack):
    """Returns (outputs [B,T,H], final states)."""

    @classmethod
    def initialize(cls, rnn_type, nlayers, hidden_size, embedding=None, input_shape=None):
        return cls(createRNN(rnn_type, nlayers, hidden_size), embedding, input_shape)

    def forward(self, x):
        return self.run(self._embed(x))


class RNNDecoder(_RNNStack):
    @classmethod
    def initialize(cls, rnn_type, nlayers, hidden_size, embedding=None, input_shape=None):
        return cls(createRNN(rnn_type, nlayers, hidden_size), embedding, input_shape)

    def forward(self, x, state):
        return self.run(self._embed(x), state)


class Bridge(nn.Module):
    """Maps every encoder state tensor to the decoder's hidden size: dense | densenonlinear (tanh) | a custom module."""

    def __init__(self, bridge_type, decoder_hidden_size, bridge=None):
        super().__init__()
        self.bridge_type, self.decoder_hidden_size, self.custom = bridge_type, decoder_hidden_size, bridge
        self.lin = None

    @classmethod
    def initialize(cls, bridge_type, decoder_hidden_size):
        if bridge_type.lower() not in ("dense", "densenonlinear"):
            raise ValueError("Only support dense|densenonlinear as bridge_type")
        return cls(bridge_type.lower(), int(decoder_hidden_size))

    @classmethod
    def initialize_from_keras_layer(cls, bridge):
        return cls("customized", 0, bridge)

    def forward(self, states):
        ts = list(states) if isinstance(states, (tuple, list)) else [states]
        if self.custom is not None:
            out = [self.custom(t) for t in ts]
        else:
            if self.lin is None:
                self.lin = nn.ModuleList([nn.Linear(t.shape[-1], self.decoder_hidden_size) for t in ts]).to(ts[0].device)
            out = [l(t) for l, t in zip(self.lin, ts)]
            if self.bridge_type == "densenonlinear":
                out = [torch.tanh(t) for t in out]
        return tuple(out) if isinstance(states, (tuple, list)) else out[0]


class Seq2seq(ZooModel):
    def __init__(self, encoder, decoder, input_shape, output_shape, bridge=None, generator=None, bigdl_type="float"):
        super().__init__()
        self.encoder, self.decoder, self.bridge, self.generator = encoder, decoder, bridge, generator
        self.input_shape, self.output_shape = list(input_shape), list(output_shape)
        if bridge is not None and bridge.custom is None:     # materialise the bridge's Linear layers now (before training)
            n_states = 2 if isinstance(encoder.rnn, nn.LSTM) else 1
            bridge.lin = nn.ModuleList([nn.Linear(encoder.hidden_size, bridge.decoder_hidden_size) for _ in range(n_states)])

    def forward(self, enc_in, dec_in):
        _, state = self.encoder(enc_in)
        if self.bridge is not None:
            state = self.bridge(state)
        out, _ = self.decoder(dec_in, state)
        return self.generator(out) if self.generator is not None else out

    def fit(self, x, y=None, batch_size=32, nb_epoch=10, validation_data=None, **kw):
        """x = [encoder input, decoder input], y = target (reference: a TextSet/RDD of such samples)."""
        return super().fit(list(x) if isinstance(x, (tuple, list)) else x, y, batch_size=batch_size, nb_epoch=nb_epoch, validation_data=validation_data)

    @torch.no_grad()
    def infer(self, input, start_sign, max_seq_len=30, stop_sign=None, build_output=None):
        """Greedy decoding: feed `start_sign`, append each step's output as the next input until `stop_sign`
        (or `max_seq_len`).  `build_output` maps a decoder output row to the next input row (default: identity)."""
        self.eval()
        dev = next(self.parameters()).device
        x = torch.as_tensor(np.asarray(input)).to(dev)
        if x.dim() == len(self.input_shape):
            x = x.unsqueeze(0)
        _, state = self.encoder(x)
        if self.bridge is not None:
            state = self.bridge(state)
        cur = torch.as_tensor(np.asarray(start_sign)).to(dev)
        if cur.dim() == len(self.output_shape) - 1:
            cur = cur.unsqueeze(0)
        cur = cur.unsqueeze(0) if cur.dim() == len(self.output_shape) else cur
        outs = []
        stop = None if stop_sign is None else torch.as_tensor(np.asarray(stop_sign)).to(dev).float().reshape(-1)
        for _ in range(max_seq_len):
            out, state = self.decoder(cur, state)
            y = self.generator(out) if self.generator is not None else out
            step = y[:, -1:]
            nxt = build_output(step) if build_output is not None else step
            outs.append(nxt)
            if stop is not None and torch.allclose(nxt.float().reshape(-1), stop, atol=1e-6):
                break
            cur = nxt
        return torch.cat(outs, dim=1).cpu().numpy()
