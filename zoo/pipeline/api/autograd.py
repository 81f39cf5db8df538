"""autograd: symbolic `Variable` algebra, math ops, `Lambda`, `Parameter`, `Constant`, `CustomLoss`.

Reference: pyzoo/zoo/pipeline/api/autograd.py:32-568 and
zoo/src/main/scala/.../pipeline/api/autograd/{math,Lambda,CustomLoss,KerasParameter}.scala.
A `Variable` is a node of a small DAG: a placeholder (`Input`), or `(module, parents)`.
`Model(input, output)` compiles the DAG into a `GraphModule` (a ``torch.nn.Module`` that evaluates the
nodes in topological order; a layer used twice shares its weights).  Differentiation is torch
autograd, so every op here is just its forward.  As in the reference, `axis` arguments count the batch
dimension as axis 0 and shapes reported by `get_output_shape` exclude it... include it as None.
"""
from __future__ import annotations

import math as _math

import numpy as np
import torch
import torch.nn as nn

from .keras.layers.base import ZooKerasLayer, _norm_shape, shape_of
from .keras.objectives import LossFunction


class _Op(nn.Module):
    """A parameter-free node: fn(*inputs)."""

    def __init__(self, fn, label="op"):
        super().__init__()
        self.fn, self.label = fn, label

    def forward(self, *xs):
        return self.fn(*xs)

    def extra_repr(self):
        return self.label


def _dry_shape(module, parent_shapes, dtypes=None):
    xs = [torch.zeros((2,) + tuple(s), dtype=(dtypes[i] if dtypes else torch.float32)) for i, s in enumerate(parent_shapes)]
    was = module.training
    module.eval()
    try:
        with torch.no_grad():
            y = module(*xs)
    finally:
        module.train(was)
    return shape_of(y)


class Variable:
    def __init__(self, input_shape=None, node=None, name=None, dtype=torch.float32):
        self.name = name
        self.module, self.parents = node if node is not None else (None, [])
        self.shape_ = _norm_shape(input_shape)
        self.dtype = dtype
        self.index = None     # for multi-output modules: which output this variable is

    # ---- construction ----------------------------------------------------------------------
    @classmethod
    def from_layer(cls, layer, parents):
        shapes = [p.shape_ for p in parents]
        if len(shapes) == 1 and isinstance(shapes[0], list):      # one multi-output variable (a table) feeding a table layer
            shapes = shapes[0]
        if isinstance(layer, ZooKerasLayer):
            out = layer.ensure_built(shapes if (len(shapes) > 1 or layer.multi_input) else shapes[0])
        else:
            out = _dry_shape(layer, shapes, [p.dtype for p in parents])
        return cls(out, (layer, list(parents)))

    @classmethod
    def from_op(cls, fn, parents, label="op"):
        parents = [p if isinstance(p, Variable) else Constant(p) for p in parents]
        op = _Op(fn, label)
        return cls(_dry_shape(op, [p.shape_ for p in parents], [p.dtype for p in parents]), (op, parents))

    # ---- introspection ---------------------------------------------------------------------
    def get_input_shape(self):
        return [(None,) + tuple(p.shape_) for p in self.parents] if self.parents else self.get_output_shape()

    def get_output_shape(self):
        return (None,) + tuple(self.shape_)

    @property
    def shape(self):
        return self.get_output_shape()

    def set_name(self, name):
        self.name = name
        return self

    # ---- operators (autograd.py:281-366) ------------------------------------------------------
    def _bin(self, other, fn, label, reverse=False):
        if isinstance(other, Variable):
            ps = [other, self] if reverse else [self, other]
            return Variable.from_op(fn, ps, label)
        c = float(other)
        return Variable.from_op((lambda a: fn(c, a)) if reverse else (lambda a: fn(a, c)), [self], label)

    def add(self, var):
        return self + var

    def sub(self, var):
        return self - var

    def __add__(self, o):
        return self._bin(o, lambda a, b: a + b, "add")

    __radd__ = __add__

    def __sub__(self, o):
        return self._bin(o, lambda a, b: a - b, "sub")

    def __rsub__(self, o):
        return self._bin(o, lambda a, b: a - b, "rsub", reverse=True)

    def __mul__(self, o):
        return self._bin(o, lambda a, b: a * b, "mul")

    __rmul__ = __mul__

    def __truediv__(self, o):
        return self._bin(o, lambda a, b: a / b, "div")

    __div__ = __truediv__

    def __rtruediv__(self, o):
        return self._bin(o, lambda a, b: a / b, "rdiv", reverse=True)

    __rdiv__ = __rtruediv__

    def __neg__(self):
        return neg(self)

    def slice(self, dim, start_index, length):
        """Narrow along `dim` (batch is dim 0): `length` elements from `start_index` (-1 = to the end)."""
        def f(x):
            n = x.shape[dim] - start_index if length == -1 else length
            return x.narrow(dim, start_index, n)

        return Variable.from_op(f, [self], "slice")

    def index_select(self, dim, index):
        """Select one index along `dim` and drop that dimension (autograd.py:340-356)."""
        return Variable.from_op(lambda x: x.select(dim, index), [self], "index_select")

    def squeeze(self, dim=None):
        return Variable.from_op((lambda x: x.squeeze()) if dim is None else (lambda x: x.squeeze(dim)), [self], "squeeze")

    def __getitem__(self, idx):
        return Variable.from_op(lambda x: x[idx], [self], "getitem")


# --------------------------------------------------------------------------- math ops --
def _unary(fn, label):
    def op(x):
        return Variable.from_op(fn, [x], label)

    op.__name__ = label
    return op


abs = _unary(torch.abs, "abs")  # noqa: A001
square = _unary(torch.square, "square")
sqrt = _unary(torch.sqrt, "sqrt")
exp = _unary(torch.exp, "exp")
log = _unary(torch.log, "log")
neg = _unary(torch.neg, "neg")
softsign = _unary(torch.nn.functional.softsign, "softsign")
softplus = _unary(torch.nn.functional.softplus, "softplus")
erf = _unary(torch.erf, "erf")
contiguous = _unary(lambda x: x.contiguous(), "contiguous")


def mean(x, axis=0, keepDims=False):
    return Variable.from_op(lambda t: t.mean(axis, keepdim=keepDims), [x], "mean")


def sum(x, axis=0, keepDims=False):  # noqa: A001
    return Variable.from_op(lambda t: t.sum(axis, keepdim=keepDims), [x], "sum")


def clip(x, min, max):  # noqa: A002
    return Variable.from_op(lambda t: t.clamp(float(min), float(max)), [x], "clip")


def pow(x, a):  # noqa: A001
    return Variable.from_op(lambda t: t.pow(float(a)), [x], "pow")


def maximum(x, y):
    if isinstance(y, Variable):
        return Variable.from_op(torch.maximum, [x, y], "maximum")
    return Variable.from_op(lambda t: t.clamp_min(float(y)), [x], "maximum")


def epsilon():
    return 1e-7


def expand_dims(x, axis):
    return Variable.from_op(lambda t: t.unsqueeze(axis), [x], "expand_dims")


def stack(inputs, axis=1):
    return Variable.from_op(lambda *ts: torch.stack(ts, axis), list(inputs), "stack")


def l2_normalize(x, axis):
    return Variable.from_op(lambda t: t / t.pow(2).sum(axis, keepdim=True).clamp_min(1e-12).sqrt(), [x], "l2_normalize")


def mm(x, y, axes=None):
    """Batched matrix multiply over the last two dims; `axes=[a, b]` contracts dim a of x with dim b of y
    (math.scala `mm`: 2 <= dims <= 4)."""
    def f(a, b):
        if axes is not None:
            ax, bx = axes
            if ax == a.dim() - 2:
                a = a.transpose(-1, -2)
            if bx == b.dim() - 1:
                b = b.transpose(-1, -2)
        return torch.matmul(a, b)

    return Variable.from_op(f, [x, y], "mm")


def batch_dot(x, y, axes=1, normalize=False):
    """Keras `batch_dot` for 2-D / 3-D inputs (autograd.py:55-77)."""
    ax = [axes, axes] if isinstance(axes, int) else list(axes)

    def f(a, b):
        if normalize:
            a = a / a.pow(2).sum(ax[0], keepdim=True).clamp_min(1e-12).sqrt()
            b = b / b.pow(2).sum(ax[1], keepdim=True).clamp_min(1e-12).sqrt()
        if a.dim() == 2 and b.dim() == 2:
            return (a * b).sum(1, keepdim=True)
        a2 = a if ax[0] == a.dim() - 1 else a.transpose(1, 2)       # contract a's last dim
        b2 = b if ax[1] == 1 else b.transpose(1, 2)                 # with b's dim 1
        return torch.bmm(a2, b2)

    return Variable.from_op(f, [x, y], "batch_dot")


# ----------------------------------------------------------------- graph execution --
class GraphModule(nn.Module):
    """Executable DAG over Variables: `GraphModule(inputs, outputs)(*tensors)`."""

    def __init__(self, inputs, outputs):
        super().__init__()
        self.inputs = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        self.single_output = not isinstance(outputs, (list, tuple))
        self.outputs = [outputs] if self.single_output else list(outputs)
        order, seen = [], set()

        def visit(v):
            if id(v) in seen:
                return
            seen.add(id(v))
            for p in v.parents:
                visit(p)
            order.append(v)

        for o in self.outputs:
            visit(o)
        self.order = order
        for v in order:
            if v.module is None and not any(v is i for i in self.inputs) and not isinstance(v, (Constant, Parameter)):
                raise ValueError("graph has a placeholder that is not listed in `input`")
        mods, ids = [], {}
        for v in order:
            if v.module is not None and id(v.module) not in ids:
                ids[id(v.module)] = len(mods)
                mods.append(v.module)
        self.mods = nn.ModuleList(mods)

    def forward(self, *xs):
        if len(xs) == 1 and isinstance(xs[0], (list, tuple)) and len(self.inputs) > 1:
            xs = tuple(xs[0])
        if len(xs) != len(self.inputs):
            raise ValueError(f"graph expects {len(self.inputs)} inputs, got {len(xs)}")
        val = {id(v): x for v, x in zip(self.inputs, xs)}
        for v in self.order:
            if id(v) in val:
                continue
            args = [val[id(p)] for p in v.parents]
            val[id(v)] = v.module(*args)
        outs = [val[id(o)] for o in self.outputs]
        return outs[0] if self.single_output else tuple(outs)

    def named_layers(self):
        return [(getattr(m, "zoo_name", type(m).__name__), m) for m in self.mods]


class _ParamModule(nn.Module):
    def __init__(self, shape, init_method, init_weight, trainable):
        super().__init__()
        w = torch.empty(tuple(shape))
        if init_weight is not None:
            w = torch.as_tensor(np.asarray(init_weight), dtype=torch.float32).reshape(tuple(shape)).clone()
        else:
            from .keras.layers.base import init_weight as _init

            _init(w, init_method or "glorot_uniform")
        self.weight = nn.Parameter(w, requires_grad=bool(trainable))

    def forward(self):
        return self.weight


class Parameter(Variable):
    """A trainable tensor usable in Variable expressions; its shape is the FULL shape (no batch dim)."""

    def __init__(self, shape, init_method=None, init_weight=None, trainable=True, name=None):
        mod = _ParamModule(shape, init_method, init_weight, trainable)
        super().__init__(tuple(shape)[1:], (mod, []), name)
        self.full_shape = tuple(shape)

    @property
    def shape(self):
        return self.full_shape

    def get_weight(self):
        return self.module.weight.detach().cpu().numpy()

    def set_weight(self, value):
        with torch.no_grad():
            self.module.weight.copy_(torch.as_tensor(np.asarray(value), dtype=torch.float32).reshape(self.full_shape))


class _ConstModule(nn.Module):
    def __init__(self, data):
        super().__init__()
        self.register_buffer("value", torch.as_tensor(np.asarray(data), dtype=torch.float32))

    def forward(self):
        return self.value


class Constant(Variable):
    def __init__(self, data, name=None, bigdl_type="float"):
        mod = _ConstModule(data)
        super().__init__(tuple(mod.value.shape)[1:] if mod.value.dim() else (), (mod, []), name)


def _dry_shape_const_aware(module, parents):
    """Parents that are Parameters/Constants contribute their full tensor, not a batched placeholder."""
    xs = []
    for p in parents:
        if isinstance(p, (Parameter, Constant)):
            xs.append(p.module())
        else:
            xs.append(torch.zeros((2,) + tuple(p.shape_), dtype=p.dtype))
    was = module.training
    module.eval()
    try:
        with torch.no_grad():
            y = module(*xs)
    finally:
        module.train(was)
    return shape_of(y)


def _from_op(fn, parents, label="op"):
    if parents and not any(isinstance(p, Variable) for p in parents) and any(torch.is_tensor(p) for p in parents):
        return fn(*parents)          # called on tensors (e.g. a loss function given to compile): evaluate eagerly, torch autograd takes over
    parents = [p if isinstance(p, Variable) else Constant(p) for p in parents]
    op = _Op(fn, label)
    return Variable(_dry_shape_const_aware(op, parents), (op, parents))


Variable.from_op = classmethod(lambda cls, fn, parents, label="op": _from_op(fn, parents, label))


class Lambda(ZooKerasLayer):
    """A layer from a function of Variables: `Lambda(lambda a, b: a + b)([x, y])` (autograd.py:393-440)."""

    def __init__(self, function, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.function = function
        self.graph = None

    def build(self, input_shape):
        shapes = input_shape if isinstance(input_shape, list) else [input_shape]
        ins = [Variable(s) for s in shapes]
        out = self.function(*ins)
        self.graph = GraphModule(ins, out)
        self.multi_input = len(shapes) > 1

    def call(self, x):
        return self.graph(*x) if isinstance(x, list) else self.graph(x)

    def create(self, input_shapes):
        self.ensure_built(input_shapes)
        return self


class LambdaLayer(Lambda):
    def __init__(self, input_vars, out_var, input_shape=None, **kw):
        ZooKerasLayer.__init__(self, input_shape, **kw)
        self.graph = GraphModule(input_vars, out_var)
        self.built = True
        self.multi_input = len(self.graph.inputs) > 1
        self.output_shape_ = out_var.shape_


class CustomLoss(LossFunction):
    """Loss from a function of (y_true, y_pred) Variables; the result is averaged over the batch
    (autograd.py:510-568, CustomLoss.scala)."""

    def __init__(self, loss_func, y_pred_shape, y_true_shape=None):
        super().__init__()
        y_true_shape = y_pred_shape if y_true_shape is None else y_true_shape
        yt, yp = Variable(y_true_shape), Variable(y_pred_shape)
        self.graph = GraphModule([yt, yp], loss_func(yt, yp))

    def forward(self, y_pred, y_true):
        out = self.graph(y_true.float().reshape(y_true.shape[0], *self.graph.inputs[0].shape_), y_pred.float())
        return out.mean()

    def backward(self, y_true, y_pred):
        yp = torch.as_tensor(np.asarray(y_pred), dtype=torch.float32).requires_grad_(True)
        self.forward(yp, torch.as_tensor(np.asarray(y_true), dtype=torch.float32)).backward()
        return yp.grad.numpy()


VariableOperator = Variable
