"""Recurrent layers (reference: layers/recurrent.py; Scala LSTM.scala, GRU.scala, SimpleRNN.scala, Recurrent.scala).
Input (batch, steps, dim) -> (batch, output_dim) or (batch, steps, output_dim) with return_sequences.
LSTM / GRU run on the persistent kernels K6 on the GPU; the zoo default `inner_activation="hard_sigmoid"`
is supported natively by those kernels."""
from __future__ import annotations

from .....ops import rnn as _rnn
from .base import ZooKerasLayer


class _RecurrentLayer(ZooKerasLayer):
    impl = None

    def __init__(self, output_dim, activation="tanh", inner_activation="hard_sigmoid", return_sequences=False, go_backwards=False,
                 W_regularizer=None, U_regularizer=None, b_regularizer=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if activation != "tanh":
            raise ValueError(f"{type(self).__name__}: only activation='tanh' is supported (got {activation})")
        self.output_dim, self.inner_activation = int(output_dim), inner_activation
        self.return_sequences, self.go_backwards = return_sequences, go_backwards

    def build(self, input_shape):
        self.cell = self.impl(input_shape[-1], self.output_dim, go_backwards=self.go_backwards, inner_activation=self.inner_activation)

    def call(self, x):
        return self.cell(x, return_sequences=self.return_sequences)


class LSTM(_RecurrentLayer):
    impl = _rnn.LSTM


class GRU(_RecurrentLayer):
    impl = _rnn.GRU


class SimpleRNN(ZooKerasLayer):
    def __init__(self, output_dim, activation="tanh", return_sequences=False, go_backwards=False, W_regularizer=None,
                 U_regularizer=None, b_regularizer=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.output_dim, self.activation = int(output_dim), activation
        self.return_sequences, self.go_backwards = return_sequences, go_backwards

    def build(self, input_shape):
        self.cell = _rnn.SimpleRNN(input_shape[-1], self.output_dim, self.activation, self.go_backwards)

    def call(self, x):
        return self.cell(x, return_sequences=self.return_sequences)
