"""Base of the keras-style layers: lazy build from an input shape, shape inference, symbolic calls.

Reference: `ZooKerasLayer` / `KerasLayer` (pyzoo/zoo/pipeline/api/keras/engine/topology.py,
zoo/src/main/scala/.../pipeline/api/keras/layers/utils/KerasUtils.scala).  A layer is a
``torch.nn.Module`` whose parameters are created by `build(input_shape)` the moment the input shape is
known: from its own `input_shape=` argument, from the previous layer in a `Sequential`, from the
`Variable` it is called on (functional API), or from the first tensor it sees.  Shapes exclude the batch
dimension, as in the reference.  `compute_output_shape` defaults to a dry run on zeros, so a layer only
has to say how to build and how to compute.
"""
from __future__ import annotations

import itertools
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

_counter = itertools.count(1)


def _norm_shape(s):
    if s is None:
        return None
    if isinstance(s, (list, tuple)) and s and isinstance(s[0], (list, tuple)):
        return [tuple(int(d) for d in x) for x in s]
    return tuple(int(d) for d in s)


def shape_of(x):
    if isinstance(x, (list, tuple)):
        return [tuple(t.shape[1:]) for t in x]
    return tuple(x.shape[1:])


# ------------------------------------------------------------------ activations --
def hard_sigmoid(x):
    return torch.clamp(0.2 * x + 0.5, 0.0, 1.0)


# module-level functions (not lambdas): layers keep a reference to their activation and must stay picklable for saveModel
def linear(x):
    return x


def softmax_last(x):
    return torch.softmax(x, -1)


def softmin_last(x):
    return F.softmin(x, -1)


def log_softmax_last(x):
    return F.log_softmax(x, -1)


ACTIVATIONS = {
    None: linear, "linear": linear, "relu": torch.relu, "tanh": torch.tanh, "sigmoid": torch.sigmoid,
    "hard_sigmoid": hard_sigmoid, "softmax": softmax_last, "softplus": F.softplus, "softsign": F.softsign,
    "relu6": F.relu6, "tanh_shrink": F.tanhshrink, "softmin": softmin_last, "log_sigmoid": F.logsigmoid,
    "log_softmax": log_softmax_last, "gelu": F.gelu, "elu": F.elu, "selu": F.selu, "exp": torch.exp,
}


def get_activation(name):
    if callable(name):
        return name
    if isinstance(name, str):
        name = name.lower()
    if name not in ACTIVATIONS:
        raise ValueError(f"Unsupported activation type: {name}")
    return ACTIVATIONS[name]


def init_weight(t: torch.Tensor, method="glorot_uniform"):
    """KerasUtils.getInitMethod: glorot_uniform | one | zero | uniform | normal (+ he/lecun for convenience)."""
    m = (method or "glorot_uniform").lower()
    with torch.no_grad():
        if m in ("glorot_uniform", "xavier"):
            if t.dim() >= 2:
                nn.init.xavier_uniform_(t)
            else:
                t.uniform_(-0.05, 0.05)
        elif m == "glorot_normal":
            nn.init.xavier_normal_(t)
        elif m in ("one", "ones"):
            t.fill_(1.0)
        elif m in ("zero", "zeros"):
            t.zero_()
        elif m == "uniform":
            t.uniform_(-0.05, 0.05)
        elif m == "normal":
            t.normal_(0.0, 0.05)
        elif m in ("he_normal", "msra"):
            nn.init.kaiming_normal_(t)
        elif m == "he_uniform":
            nn.init.kaiming_uniform_(t)
        elif m == "lecun_uniform":
            fan_in = t.shape[1] if t.dim() >= 2 else t.numel()
            b = math.sqrt(3.0 / max(fan_in, 1))
            t.uniform_(-b, b)
        else:
            raise ValueError(f"Unsupported init method: {method}")
    return t


class ZooKerasLayer(nn.Module):
    input_dtype = torch.float32
    multi_input = False

    def __init__(self, input_shape=None, name=None, bigdl_type="float", **kwargs):
        super().__init__()
        if kwargs:
            raise TypeError(f"{type(self).__name__}: unexpected arguments {sorted(kwargs)}")
        self.zoo_name = name or f"{type(self).__name__.lower()}{next(_counter)}"
        self._declared_input_shape = _norm_shape(input_shape)
        self.built = False
        self.input_shape_ = None
        self.output_shape_ = None

    # -- to override ------------------------------------------------------------------------
    def build(self, input_shape):
        pass

    def call(self, x):  # pragma: no cover - interface
        raise NotImplementedError

    def compute_output_shape(self, input_shape):
        """Default: dry run on zeros (batch of 2) in eval mode."""
        mk = lambda s: torch.zeros((2,) + tuple(s), dtype=self.input_dtype)
        x = [mk(s) for s in input_shape] if isinstance(input_shape, list) else mk(input_shape)
        was = self.training
        self.eval()
        try:
            with torch.no_grad():
                y = self.call(x)
        finally:
            self.train(was)
        return shape_of(y)

    # -- plumbing ---------------------------------------------------------------------------
    def ensure_built(self, input_shape):
        input_shape = _norm_shape(input_shape)
        if not self.built:
            if input_shape is None:
                raise ValueError(f"{self.zoo_name}: the first layer of a model needs input_shape=")
            self.build(input_shape)
            self.built = True
            self.input_shape_ = input_shape
            self.output_shape_ = self.compute_output_shape(input_shape)
        return self.output_shape_

    build_from = ensure_built

    def forward(self, *xs):
        x = xs[0] if len(xs) == 1 else list(xs)
        if not self.built:
            self.ensure_built(shape_of(x))
            dev = (x[0] if isinstance(x, list) else x).device
            self.to(dev)
        return self.call(x)

    def __call__(self, *args, **kwargs):
        from ...autograd import Variable

        flat = args[0] if (len(args) == 1 and isinstance(args[0], (list, tuple))) else args
        if flat and all(isinstance(a, Variable) for a in flat):
            return Variable.from_layer(self, list(flat))
        return super().__call__(*args, **kwargs)

    def get_input_shape(self):
        s = self.input_shape_ if self.input_shape_ is not None else self._declared_input_shape
        if s is None:
            return None
        return [(None,) + tuple(x) for x in s] if isinstance(s, list) else (None,) + tuple(s)

    def get_output_shape(self):
        s = self.output_shape_
        if s is None:
            return None
        return [(None,) + tuple(x) for x in s] if isinstance(s, list) else (None,) + tuple(s)

    def set_name(self, name):
        self.zoo_name = name
        return self

    def name(self):
        return self.zoo_name

    def get_weights(self):
        return [p.detach().cpu().numpy() for p in self.parameters()]

    def set_weights(self, weights):
        ps = list(self.parameters())
        if len(ps) != len(weights):
            raise ValueError(f"{self.zoo_name}: expected {len(ps)} weight arrays, got {len(weights)}")
        with torch.no_grad():
            for p, w in zip(ps, weights):
                p.copy_(torch.as_tensor(w).reshape(p.shape))
        return self

    def extra_repr(self):
        return f"name={self.zoo_name}, output_shape={self.output_shape_}"


class _FnLayer(ZooKerasLayer):
    """Parameter-free layer defined by a function of the input tensor."""

    def fn(self, x):  # pragma: no cover - interface
        raise NotImplementedError

    def call(self, x):
        return self.fn(x)


class InputLayer(_FnLayer):
    def __init__(self, input_shape=None, **kw):
        super().__init__(input_shape, **kw)

    def fn(self, x):
        return x


def Input(shape=None, name=None):
    """Symbolic input of the functional API (topology.py `Input`)."""
    from ...autograd import Variable

    return Variable(input_shape=shape, name=name)
