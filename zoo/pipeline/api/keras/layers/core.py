"""Core layers (reference: pyzoo/zoo/pipeline/api/keras/layers/core.py; Scala Dense.scala, SparseDense.scala,
MaxoutDense.scala, Highway.scala, Merge.scala, ...)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .....ops.dense import FusedDense
from .base import ZooKerasLayer, _FnLayer, get_activation, init_weight


class Masking(_FnLayer):
    """Zero every timestep whose features all equal `mask_value`."""

    def __init__(self, mask_value=0.0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.mask_value = float(mask_value)

    def fn(self, x):
        keep = (x != self.mask_value).any(-1, keepdim=True)
        return x * keep.to(x.dtype)


class Dropout(_FnLayer):
    def __init__(self, p, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.p = float(p)

    def fn(self, x):
        return torch.nn.functional.dropout(x, self.p, self.training)


class _SpatialDropout(_FnLayer):
    def __init__(self, p=0.5, dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.p, self.dim_ordering = float(p), dim_ordering

    def fn(self, x):
        if not self.training or self.p == 0:
            return x
        shape = list(x.shape)
        if x.dim() == 3:                       # 1D: (batch, steps, dim) -> drop whole feature channels
            shape[1] = 1
        elif self.dim_ordering == "th":
            shape[2:] = [1] * (x.dim() - 2)
        else:
            shape[1:-1] = [1] * (x.dim() - 2)
        mask = torch.bernoulli(torch.full(shape, 1 - self.p, device=x.device, dtype=x.dtype)) / (1 - self.p)
        return x * mask


class SpatialDropout1D(_SpatialDropout):
    def __init__(self, p=0.5, input_shape=None, **kw):
        super().__init__(p, "th", input_shape, **kw)


class SpatialDropout2D(_SpatialDropout):
    pass


class SpatialDropout3D(_SpatialDropout):
    pass


class Activation(_FnLayer):
    def __init__(self, activation, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.activation = activation
        self._fn = get_activation(activation)
        self.final_activation_kind = activation if activation in ("softmax", "sigmoid") else None

    def fn(self, x):
        return self._fn(x)

    def strip_final_activation(self, x):
        return x


class Reshape(_FnLayer):
    """`target_shape` may hold one -1 (inferred)."""

    def __init__(self, target_shape, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.target_shape = tuple(int(d) for d in target_shape)

    def fn(self, x):
        return x.reshape(x.shape[0], *self.target_shape)


class Permute(_FnLayer):
    """`dims` are 1-based and exclude the batch dimension: Permute((2, 1))."""

    def __init__(self, dims, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dims = tuple(int(d) for d in dims)

    def fn(self, x):
        return x.permute(0, *self.dims).contiguous()


class Flatten(_FnLayer):
    def fn(self, x):
        return x.reshape(x.shape[0], -1)


class RepeatVector(_FnLayer):
    def __init__(self, n, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.n = int(n)

    def fn(self, x):
        return x.unsqueeze(1).expand(-1, self.n, -1)


class Dense(ZooKerasLayer):
    """Fully connected layer on the last dimension; on the GPU the GEMM + bias + ReLU run on the tcgen05
    kernel (zoo.ops.dense).  A trailing softmax / sigmoid can be peeled off for the fused loss kernels."""

    def __init__(self, output_dim, init="glorot_uniform", activation=None, W_regularizer=None, b_regularizer=None, bias=True,
                 input_dim=None, input_shape=None, **kw):
        if input_dim is not None and input_shape is None:
            input_shape = (int(input_dim),)
        super().__init__(input_shape, **kw)
        self.output_dim, self.init, self.activation, self.use_bias = int(output_dim), init, activation, bias
        self.W_regularizer, self.b_regularizer = W_regularizer, b_regularizer
        self.final_activation_kind = activation if activation in ("softmax", "sigmoid") else None

    def build(self, input_shape):
        fused = self.activation if self.activation in (None, "linear", "relu") else None
        self.dense = FusedDense(input_shape[-1], self.output_dim, activation=fused, bias=self.use_bias)
        init_weight(self.dense.weight, self.init)
        self._post = None if fused == self.activation or self.activation in (None, "linear") else get_activation(self.activation)
        self._zoo_regularizers = [(r, t) for r, t in ((self.W_regularizer, self.dense.weight), (self.b_regularizer, self.dense.bias))
                                  if r is not None and t is not None]

    def strip_final_activation(self, x):
        return self.dense(x)

    def call(self, x):
        y = self.dense(x)
        return y if self._post is None else self._post(y.float())


class SparseDense(Dense):
    """Dense whose input is sparse (reference SparseDense.scala: no gradient w.r.t. the input).  Inputs may be a
    torch sparse tensor or a dense tensor; `backward_start/length` are accepted for API parity."""

    def __init__(self, output_dim, init="glorot_uniform", activation=None, W_regularizer=None, b_regularizer=None,
                 backward_start=-1, backward_length=-1, init_weight=None, init_bias=None, init_grad_weight=None,
                 init_grad_bias=None, bias=True, input_shape=None, **kw):
        super().__init__(output_dim, init, activation, W_regularizer, b_regularizer, bias, None, input_shape, **kw)
        self._init_w, self._init_b = init_weight, init_bias

    def build(self, input_shape):
        super().build(input_shape)
        with torch.no_grad():
            if self._init_w is not None:
                self.dense.weight.copy_(torch.as_tensor(self._init_w, dtype=torch.float32).reshape(self.dense.weight.shape))
            if self._init_b is not None and self.dense.bias is not None:
                self.dense.bias.copy_(torch.as_tensor(self._init_b, dtype=torch.float32).reshape(-1))

    def call(self, x):
        if x.is_sparse:
            y = torch.sparse.mm(x.float(), self.dense.weight.t())
            if self.dense.bias is not None:
                y = y + self.dense.bias
            return get_activation(self.activation)(y)
        return super().call(x.detach())


class MaxoutDense(ZooKerasLayer):
    def __init__(self, output_dim, nb_feature=4, W_regularizer=None, b_regularizer=None, bias=True, input_dim=None,
                 input_shape=None, **kw):
        if input_dim is not None and input_shape is None:
            input_shape = (int(input_dim),)
        super().__init__(input_shape, **kw)
        self.output_dim, self.nb_feature, self.use_bias = int(output_dim), int(nb_feature), bias

    def build(self, input_shape):
        self.lin = nn.Linear(input_shape[-1], self.output_dim * self.nb_feature, bias=self.use_bias)

    def call(self, x):
        return self.lin(x.float()).reshape(*x.shape[:-1], self.nb_feature, self.output_dim).max(-2).values


class Highway(ZooKerasLayer):
    """y = T * act(W x + b) + (1 - T) * x with T = sigmoid(W_T x + b_T)."""

    def __init__(self, activation=None, W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.activation, self.use_bias = activation, bias

    def build(self, input_shape):
        d = input_shape[-1]
        self.h = nn.Linear(d, d, bias=self.use_bias)
        self.t = nn.Linear(d, d, bias=self.use_bias)

    def call(self, x):
        x = x.float()
        t = torch.sigmoid(self.t(x))
        return t * get_activation(self.activation)(self.h(x)) + (1 - t) * x


class GetShape(_FnLayer):
    def fn(self, x):
        return torch.tensor(list(x.shape), dtype=torch.float32, device=x.device)

    def compute_output_shape(self, input_shape):
        return (len(input_shape) + 1,)


class Max(_FnLayer):
    """Max (values, or indices with return_value=False) over `dim` (1-based... 0-based excluding batch in the
    python API: dim counts from the first non-batch dimension)."""

    def __init__(self, dim, num_input_dims=-1, return_value=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dim, self.return_value = int(dim), return_value

    def fn(self, x):
        d = self.dim + 1 if self.dim >= 0 else self.dim
        r = x.max(d)
        return r.values if self.return_value else r.indices.to(x.dtype)


class ExpandDim(_FnLayer):
    def __init__(self, dim, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dim = int(dim)

    def fn(self, x):
        return x.unsqueeze(self.dim)


class Merge(ZooKerasLayer):
    """Merge a list of inputs: sum | mul | concat | ave | cos | dot | max | sub | div | min (Merge.scala)."""
    multi_input = True

    def __init__(self, layers=None, mode="sum", concat_axis=-1, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.mode, self.concat_axis = mode.lower(), int(concat_axis)
        if self.mode not in ("sum", "mul", "concat", "ave", "cos", "dot", "max", "sub", "div", "min"):
            raise ValueError(f"Invalid merge mode: {mode}")
        self.branches = nn.ModuleList(layers or [])

    def build(self, input_shape):
        for b, s in zip(self.branches, input_shape):
            if isinstance(b, ZooKerasLayer):
                b.ensure_built(s)

    def call(self, xs):
        if len(self.branches):
            xs = [b(x) for b, x in zip(self.branches, xs)]
        m = self.mode
        if m == "concat":
            ax = self.concat_axis if self.concat_axis < 0 else self.concat_axis
            dt = xs[0].dtype
            return torch.cat([x.to(dt) for x in xs], dim=ax)
        xs = [x.float() for x in xs]
        if m in ("sub", "div", "cos", "dot") and len(xs) != 2:
            raise ValueError(f"merge mode {m} takes exactly two inputs")
        if m == "sum":
            return torch.stack(xs).sum(0)
        if m == "mul":
            out = xs[0]
            for x in xs[1:]:
                out = out * x
            return out
        if m == "ave":
            return torch.stack(xs).mean(0)
        if m == "max":
            return torch.stack(xs).max(0).values
        if m == "min":
            return torch.stack(xs).min(0).values
        if m == "sub":
            return xs[0] - xs[1]
        if m == "div":
            return xs[0] / xs[1]
        a, b = xs[0].reshape(xs[0].shape[0], -1), xs[1].reshape(xs[1].shape[0], -1)
        if m == "dot":
            return (a * b).sum(1, keepdim=True)
        return torch.nn.functional.cosine_similarity(a, b, dim=1).reshape(-1, 1, 1)


def merge(inputs, mode="sum", concat_axis=-1, name=None):
    """Functional merge of Variables."""
    return Merge(mode=mode, concat_axis=concat_axis, name=name)(list(inputs))
