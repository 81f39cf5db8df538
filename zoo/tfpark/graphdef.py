"""TensorFlow frozen graphs without TensorFlow: read (and write) `GraphDef` protobufs and run them as a torch module.

The reference executes frozen graphs through libtensorflow (`TFNet`, zoo/.../pipeline/api/net/TFNet.scala:56-716; Python side
pyzoo/zoo/tfpark/tfnet.py:52-302 and `export_tf`, pyzoo/zoo/util/tf.py:50-238).  There is no TF runtime here, so a frozen
inference graph is IMPORTED: the GraphDef is decoded with :mod:`zoo.util.protowire`, constants become buffers of an
`nn.Module`, and the nodes are evaluated by torch ops (NHWC tensors are kept in TensorFlow's layout; convolutions and
poolings permute around the torch call).  Covered: the ops of feed-forward / CNN inference graphs (see `_OPS`); anything else
raises `NotImplementedError` naming the node.  `write_graphdef` is the inverse, used by `export_tf` and by the tests.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ..util import protowire as pw

# ---- schemas (tensorflow/core/framework/{graph,node_def,attr_value,tensor,tensor_shape,types}.proto) ------------------------------
DIM = {1: ("size", "varint"), 2: ("name", "string")}
SHAPE = {2: ("dim", "msg", DIM), 3: ("unknown_rank", "bool")}
TENSOR = {1: ("dtype", "enum"), 2: ("tensor_shape", "msg", SHAPE), 4: ("tensor_content", "bytes"), 5: ("float_val", "packed_float"),
          6: ("double_val", "packed_double"), 7: ("int_val", "packed_varint"), 8: ("string_val", "bytes"), 10: ("int64_val", "packed_varint"),
          11: ("bool_val", "packed_varint"), 13: ("half_val", "packed_varint")}
LIST = {2: ("s", "bytes"), 3: ("i", "packed_varint"), 4: ("f", "packed_float"), 5: ("b", "packed_varint"), 6: ("type", "packed_varint"),
        7: ("shape", "msg", SHAPE), 8: ("tensor", "msg", TENSOR)}
ATTR = {1: ("list", "msg", LIST), 2: ("s", "bytes"), 3: ("i", "varint"), 4: ("f", "float"), 5: ("b", "bool"), 6: ("type", "enum"),
        7: ("shape", "msg", SHAPE), 8: ("tensor", "msg", TENSOR)}
ATTR_ENTRY = {1: ("key", "string"), 2: ("value", "msg", ATTR)}
NODE = {1: ("name", "string"), 2: ("op", "string"), 3: ("input", "string"), 4: ("device", "string"), 5: ("attr", "msg", ATTR_ENTRY)}
GRAPH = {1: ("node", "msg", NODE)}

DT = {1: np.float32, 2: np.float64, 3: np.int32, 4: np.uint8, 5: np.int16, 6: np.int8, 9: np.int64, 10: np.bool_, 19: np.float16}
DT_OF = {np.dtype(v): k for k, v in DT.items()}


def _tensor_to_numpy(t):
    dtype = DT.get(pw.one(t, "dtype", 1))
    if dtype is None:
        raise NotImplementedError(f"GraphDef tensor dtype {pw.one(t, 'dtype')} is not supported")
    shape = [int(pw.one(d, "size", 0)) for d in pw.one(t, "tensor_shape", {}).get("dim", [])]
    n = int(np.prod(shape)) if shape else 1
    if t.get("tensor_content"):
        return np.frombuffer(t["tensor_content"][-1], dtype=dtype).reshape(shape).copy()
    for key in ("float_val", "double_val", "int_val", "int64_val", "bool_val"):
        if t.get(key):
            vals = np.concatenate([np.atleast_1d(np.asarray(v)) for v in t[key]]).astype(dtype)
            if vals.size == 1 and n > 1:
                vals = np.full(n, vals[0], dtype=dtype)       # splat encoding
            return vals.reshape(shape)
    return np.zeros(shape, dtype=dtype)


class Node:
    __slots__ = ("name", "op", "inputs", "attr")

    def __init__(self, name, op, inputs, attr):
        self.name, self.op, self.inputs, self.attr = name, op, inputs, attr

    def a(self, key, default=None):
        v = self.attr.get(key)
        if v is None:
            return default
        for k in ("i", "f", "b", "type"):
            if k in v:
                return v[k][-1]
        if "s" in v:
            return v["s"][-1].decode()
        if "list" in v:
            lst = v["list"][-1]
            for k in ("i", "f", "b", "type"):
                if k in lst:
                    return list(lst[k])
            if "s" in lst:
                return [x.decode() for x in lst["s"]]
            return []
        if "shape" in v:
            return [int(pw.one(d, "size", -1)) for d in v["shape"][-1].get("dim", [])]
        if "tensor" in v:
            return _tensor_to_numpy(v["tensor"][-1])
        return default


def read_graphdef(path_or_bytes):
    data = path_or_bytes
    if not isinstance(data, (bytes, bytearray, memoryview)):
        with open(data, "rb") as f:
            data = f.read()
    g = pw.decode(data, GRAPH)
    nodes = []
    for n in g.get("node", []):
        attr = {pw.one(e, "key"): pw.one(e, "value", {}) for e in n.get("attr", [])}
        nodes.append(Node(pw.one(n, "name"), pw.one(n, "op"), [str(i) for i in n.get("input", [])], attr))
    return nodes


def _attr_msg(v):
    if isinstance(v, bool):
        return {"b": v}
    if isinstance(v, int):
        return {"i": v}
    if isinstance(v, float):
        return {"f": v}
    if isinstance(v, str):
        return {"s": v.encode()}
    if isinstance(v, np.dtype) or (isinstance(v, type) and issubclass(v, np.generic)):
        return {"type": DT_OF[np.dtype(v)]}
    if isinstance(v, np.ndarray):
        return {"tensor": {"dtype": DT_OF[v.dtype], "tensor_shape": {"dim": [{"size": int(s)} for s in v.shape]}, "tensor_content": v.tobytes()}}
    if isinstance(v, (list, tuple)):
        if all(isinstance(x, int) for x in v):
            return {"list": {"i": list(v)}}
        if all(isinstance(x, float) for x in v):
            return {"list": {"f": list(v)}}
    raise TypeError(f"cannot encode attribute value {v!r}")


def write_graphdef(path, nodes):
    """nodes: [(name, op, [inputs], {attr: python value})] -> binary GraphDef at `path` (returns the bytes)."""
    msg = {"node": [{"name": n, "op": op, "input": list(ins), "attr": [{"key": k, "value": _attr_msg(v)} for k, v in attrs.items()]}
                    for n, op, ins, attrs in nodes]}
    data = pw.encode(msg, GRAPH)
    if path is not None:
        with open(path, "wb") as f:
            f.write(data)
    return data


# ---- op implementations: fn(node, *input tensors) -> tensor or tuple ------------------------------------------------------------
def _nhwc(node):
    return node.a("data_format", "NHWC") in ("NHWC", None)


def _same_pad(x, k, s, d=(1, 1)):
    """TensorFlow SAME padding for an NCHW tensor (may be asymmetric: extra pixel goes to the bottom / right)."""
    H, W = x.shape[-2:]
    pads = []
    for size, kk, ss, dd in ((W, k[1], s[1], d[1]), (H, k[0], s[0], d[0])):
        out = -(-size // ss)
        total = max((out - 1) * ss + (kk - 1) * dd + 1 - size, 0)
        pads += [total // 2, total - total // 2]
    return pads


def _conv(node, x, w, depthwise=False):
    nhwc = _nhwc(node)
    st, dl = node.a("strides", [1, 1, 1, 1]), node.a("dilations", [1, 1, 1, 1])
    s = (st[1], st[2]) if nhwc else (st[2], st[3])
    d = (dl[1], dl[2]) if nhwc else (dl[2], dl[3])
    xin = x.permute(0, 3, 1, 2) if nhwc else x
    kh, kw, cin, mult = w.shape
    if depthwise:
        wt = w.permute(2, 3, 0, 1).reshape(cin * mult, 1, kh, kw)
        groups = cin
    else:
        wt = w.permute(3, 2, 0, 1)
        groups = 1
    if node.a("padding", "VALID") == "SAME":
        xin = F.pad(xin, _same_pad(xin, (kh, kw), s, d))
    y = F.conv2d(xin, wt.to(xin.dtype), None, s, 0, d, groups)
    return y.permute(0, 2, 3, 1) if nhwc else y


def _pool(node, x, kind):
    nhwc = _nhwc(node)
    ks, st = node.a("ksize"), node.a("strides")
    k = (ks[1], ks[2]) if nhwc else (ks[2], ks[3])
    s = (st[1], st[2]) if nhwc else (st[2], st[3])
    xin = x.permute(0, 3, 1, 2) if nhwc else x
    if node.a("padding", "VALID") == "SAME":
        pads = _same_pad(xin, k, s)
        if kind == "max":
            xin = F.pad(xin, pads, value=float("-inf"))
            y = F.max_pool2d(xin, k, s)
        else:   # TF averages over the VALID part of each window
            ones = F.pad(torch.ones_like(xin[:1, :1]), pads)
            y = F.avg_pool2d(F.pad(xin, pads), k, s) / F.avg_pool2d(ones, k, s)
    else:
        y = F.max_pool2d(xin, k, s) if kind == "max" else F.avg_pool2d(xin, k, s)
    return y.permute(0, 2, 3, 1) if nhwc else y


def _bias_add(node, x, b):
    if _nhwc(node) or x.dim() <= 2:
        return x + b
    return x + b.view([1, -1] + [1] * (x.dim() - 2))


def _fused_bn(node, x, scale, offset, mean, var):
    eps = node.a("epsilon", 1e-3)
    shape = [1] * (x.dim() - 1) + [-1] if _nhwc(node) else [1, -1] + [1] * (x.dim() - 2)
    y = (x - mean.view(shape)) * torch.rsqrt(var.view(shape) + eps) * scale.view(shape) + offset.view(shape)
    return (y, mean, var, mean, var, mean)


def _reduce(fn):
    def run(node, x, axes):
        ax = [int(a) for a in np.atleast_1d(axes.cpu().numpy())]
        return fn(x, dim=ax, keepdim=bool(node.a("keep_dims", False)))
    return run


def _strided_slice(node, x, begin, end, strides):
    b, e, s = (t.cpu().numpy().astype(np.int64).tolist() for t in (begin, end, strides))
    bm, em, sm = node.a("begin_mask", 0), node.a("end_mask", 0), node.a("shrink_axis_mask", 0)
    if node.a("ellipsis_mask", 0) or node.a("new_axis_mask", 0):
        raise NotImplementedError(f"StridedSlice {node.name}: ellipsis / new-axis masks are not supported")
    idx = []
    for i in range(len(b)):
        if sm >> i & 1:
            idx.append(b[i])
        else:
            idx.append(slice(None if bm >> i & 1 else b[i], None if em >> i & 1 else e[i], s[i]))
    return x[tuple(idx)]


def _squeeze(node, x):
    dims = node.a("squeeze_dims")
    if not dims:
        return x.squeeze()
    for d in sorted((d % x.dim() for d in dims), reverse=True):
        x = x.squeeze(d)
    return x


def _shape_list(t):
    return [int(v) for v in t.cpu().numpy().reshape(-1)]


_T = {1: torch.float32, 2: torch.float64, 3: torch.int32, 4: torch.uint8, 5: torch.int16, 6: torch.int8, 9: torch.int64, 10: torch.bool, 19: torch.float16}

_OPS = {
    "Identity": lambda n, x: x, "StopGradient": lambda n, x: x, "PlaceholderWithDefault": lambda n, x: x, "Snapshot": lambda n, x: x,
    "MatMul": lambda n, a, b: (a.t() if n.a("transpose_a", False) else a) @ (b.t() if n.a("transpose_b", False) else b),
    "BatchMatMul": lambda n, a, b: (a.transpose(-1, -2) if n.a("adj_x", False) else a) @ (b.transpose(-1, -2) if n.a("adj_y", False) else b),
    "BatchMatMulV2": lambda n, a, b: (a.transpose(-1, -2) if n.a("adj_x", False) else a) @ (b.transpose(-1, -2) if n.a("adj_y", False) else b),
    "BiasAdd": _bias_add,
    "Add": lambda n, a, b: a + b, "AddV2": lambda n, a, b: a + b, "Sub": lambda n, a, b: a - b, "Mul": lambda n, a, b: a * b,
    "RealDiv": lambda n, a, b: a / b, "Maximum": lambda n, a, b: torch.maximum(a, b), "Minimum": lambda n, a, b: torch.minimum(a, b),
    "AddN": lambda n, *xs: sum(xs[1:], xs[0]), "Pow": lambda n, a, b: torch.pow(a, b), "SquaredDifference": lambda n, a, b: (a - b) ** 2,
    "Relu": lambda n, x: F.relu(x), "Relu6": lambda n, x: F.relu6(x), "Sigmoid": lambda n, x: torch.sigmoid(x), "Tanh": lambda n, x: torch.tanh(x),
    "Elu": lambda n, x: F.elu(x), "Softplus": lambda n, x: F.softplus(x), "LeakyRelu": lambda n, x: F.leaky_relu(x, n.a("alpha", 0.2)),
    "Softmax": lambda n, x: torch.softmax(x, -1), "LogSoftmax": lambda n, x: torch.log_softmax(x, -1),
    "Rsqrt": lambda n, x: torch.rsqrt(x), "Sqrt": lambda n, x: torch.sqrt(x), "Square": lambda n, x: x * x, "Exp": lambda n, x: torch.exp(x),
    "Log": lambda n, x: torch.log(x), "Neg": lambda n, x: -x, "Abs": lambda n, x: torch.abs(x), "Erf": lambda n, x: torch.erf(x),
    "Conv2D": lambda n, x, w: _conv(n, x, w), "DepthwiseConv2dNative": lambda n, x, w: _conv(n, x, w, depthwise=True),
    "MaxPool": lambda n, x: _pool(n, x, "max"), "AvgPool": lambda n, x: _pool(n, x, "avg"),
    "FusedBatchNorm": _fused_bn, "FusedBatchNormV2": _fused_bn, "FusedBatchNormV3": _fused_bn,
    "Reshape": lambda n, x, s: x.reshape(_shape_list(s)), "Squeeze": lambda n, x: _squeeze(n, x),
    "ExpandDims": lambda n, x, d: x.unsqueeze(int(d)), "ConcatV2": lambda n, *xs: torch.cat(xs[:-1], int(xs[-1])),
    "Pack": lambda n, *xs: torch.stack(xs, n.a("axis", 0)), "Transpose": lambda n, x, p: x.permute(_shape_list(p)),
    "Pad": lambda n, x, p: F.pad(x, [v for pair in reversed(p.cpu().numpy().astype(int).tolist()) for v in pair]),
    "Mean": _reduce(torch.mean), "Sum": _reduce(torch.sum), "Max": _reduce(torch.amax), "Min": _reduce(torch.amin),
    "Shape": lambda n, x: torch.tensor(list(x.shape), dtype=torch.int32), "StridedSlice": _strided_slice,
    "Cast": lambda n, x: x.to(_T[n.a("DstT", 1)]), "ArgMax": lambda n, x, d: torch.argmax(x, int(d)),
    "GatherV2": lambda n, p, i, ax: torch.index_select(p, int(ax), i.reshape(-1).long()).reshape(
        list(p.shape[:int(ax)]) + list(i.shape) + list(p.shape[int(ax) + 1:])),
    "Tile": lambda n, x, r: x.repeat(*_shape_list(r)), "Fill": lambda n, s, v: torch.full(_shape_list(s), float(v)),
    "Greater": lambda n, a, b: a > b, "Less": lambda n, a, b: a < b, "Equal": lambda n, a, b: a == b,
    "Select": lambda n, c, a, b: torch.where(c, a, b), "SelectV2": lambda n, c, a, b: torch.where(c, a, b),
}


class _Dead:
    """Value of a tensor on the branch a `Switch` did not take (TensorFlow's dead tensors): every op fed by it is skipped, `Merge`
    forwards its live input.  Keras graphs pick the training or the inference form of BatchNorm / Dropout this way."""

    def __repr__(self):
        return "<dead tensor>"


_DEAD = _Dead()


def _switch(n, x, pred):
    take = bool(pred.reshape(-1)[0]) if torch.is_tensor(pred) else bool(pred)
    return (_DEAD, x) if take else (x, _DEAD)          # output 0 = false branch, output 1 = true branch


_OPS["Switch"] = _switch
_OPS["RefSwitch"] = _switch


def _string_to_number(n, x):
    """Strings cannot live in a torch tensor: a string input is a list / object array of str or bytes."""
    flat = np.asarray(x, dtype=object).reshape(-1)
    vals = [float(v.decode() if isinstance(v, (bytes, bytearray)) else v) for v in flat]
    return torch.tensor(vals, dtype=_T[n.a("out_type", 1)]).reshape(np.asarray(x, dtype=object).shape)


# the gradient ops `export_tf(generate_backward=True)` leaves in a graph (reference: TFNet backward through exported gradients,
# TFNet.scala:300-380); torch autograd differentiates the imported forward ops anyway, these let the exported nodes be evaluated
_OPS.update({
    "StringToNumber": _string_to_number,
    "BiasAddGrad": lambda n, g: g.sum(tuple(range(g.dim() - 1))) if _nhwc(n) else g.sum(tuple(d for d in range(g.dim()) if d != 1)),
    "ReluGrad": lambda n, g, features: g * (features > 0).to(g.dtype),
    "Relu6Grad": lambda n, g, features: g * ((features > 0) & (features < 6)).to(g.dtype),
    "SigmoidGrad": lambda n, y, dy: dy * y * (1 - y),
    "TanhGrad": lambda n, y, dy: dy * (1 - y * y),
    "SoftplusGrad": lambda n, g, features: g * torch.sigmoid(features),
})


class GraphDefModule(nn.Module):
    """A frozen GraphDef as a torch module: `forward(*inputs)` feeds `input_names` (placeholders) and returns `output_names`."""

    def __init__(self, nodes, input_names=None, output_names=None, trainable=None):
        """`trainable`: names of constants (frozen variables, e.g. "dense/kernel:0") that become `nn.Parameter`s instead of buffers,
        so the imported graph can be fine-tuned -- the reference trains a TFNet through the gradients its export wrote out."""
        super().__init__()
        self.nodes = {n.name: n for n in nodes}
        consumed = {i.split(":")[0].lstrip("^") for n in nodes for i in n.inputs}
        self.input_names = [self._strip(n) for n in (input_names or [n.name for n in nodes if n.op == "Placeholder"])]
        self.output_names = list(output_names or [n.name for n in nodes if n.name not in consumed and n.op not in ("NoOp", "Const", "Placeholder")])
        for name in self.output_names:
            if self._strip(name) not in self.nodes:
                raise KeyError(f"output {name!r} is not a node of the graph")
        # evaluation order: the nodes the outputs depend on, inputs first (iterative DFS: graphs can be thousands of nodes deep)
        self.order, seen, stack = [], set(self.input_names), [(self._strip(o), False) for o in reversed(self.output_names)]
        while stack:
            name, done = stack.pop()
            if done:
                self.order.append(name)
                continue
            if name in seen:
                continue
            seen.add(name)
            if name not in self.nodes:
                raise KeyError(f"node {name!r} is referenced but not defined")
            stack.append((name, True))
            for i in reversed(self.nodes[name].inputs):
                dep = self._strip(i.lstrip("^"))
                if dep not in seen and not i.startswith("^"):
                    stack.append((dep, False))

        # constants the outputs depend on become buffers (a training graph also carries string constants of its saver, unused here)
        self._const, needed, train = {}, set(self.order), {self._strip(v) for v in (trainable or [])}
        for k, n in enumerate(nodes):
            if n.op == "Const" and n.name in needed:
                arr = n.a("value")
                # (ascontiguousarray would turn a scalar into shape [1])
                t = torch.from_numpy(arr.copy(order="C")) if arr.dtype != np.float64 else torch.from_numpy(arr.astype(np.float32))
                key = f"c{k}"
                if n.name in train and t.is_floating_point():
                    self.register_parameter(key, nn.Parameter(t))
                else:
                    self.register_buffer(key, t, persistent=True)
                self._const[n.name] = key

    @staticmethod
    def _strip(name):
        return name.split(":")[0]

    def forward(self, *xs):
        if len(xs) == 1 and isinstance(xs[0], (tuple, list)) and not (len(self.input_names) == 1 and xs[0] and isinstance(xs[0][0], (str, bytes))):
            xs = tuple(xs[0])              # (a list of strings is ONE string input)
        if len(xs) != len(self.input_names):
            raise ValueError(f"expected {len(self.input_names)} inputs ({self.input_names}), got {len(xs)}")
        dev = xs[0].device if xs and torch.is_tensor(xs[0]) else None
        as_t = lambda x: x if torch.is_tensor(x) else (x if np.asarray(x).dtype.kind in "OSU" else torch.as_tensor(np.asarray(x)))   # strings stay lists
        vals = {n: as_t(x) for n, x in zip(self.input_names, xs)}

        def get(ref):
            name, _, port = ref.partition(":")
            v = vals[name]
            return v[int(port) if port else 0] if isinstance(v, tuple) else v

        for name in self.order:
            node = self.nodes[name]
            if node.op == "Const":
                vals[name] = getattr(self, self._const[name])
                continue
            if node.op == "Placeholder":
                raise ValueError(f"placeholder {name!r} was not fed (inputs: {self.input_names})")
            args = [get(i) for i in node.inputs if not i.startswith("^")]
            if node.op in ("Merge", "RefMerge"):
                live = [(k, a) for k, a in enumerate(args) if a is not _DEAD]
                vals[name] = (live[0][1], torch.tensor(live[0][0], dtype=torch.int32)) if live else (_DEAD, _DEAD)
                continue
            if any(a is _DEAD for a in args):
                vals[name] = (_DEAD, _DEAD) if node.op in ("Switch", "RefSwitch") else _DEAD
                continue
            fn = _OPS.get(node.op)
            if fn is None:
                raise NotImplementedError(f"TensorFlow op {node.op!r} (node {name!r}) is not supported by the GraphDef importer")
            if dev is not None:       # tensors created on the host by shape arithmetic follow the data
                args = [a.to(dev) if torch.is_tensor(a) and a.device != dev and a.is_floating_point() else a for a in args]
            vals[name] = fn(node, *args)
        outs = [get(o) for o in self.output_names]
        if any(o is _DEAD for o in outs):
            raise ValueError("an output sits on a branch the graph's Switch ops did not take")
        return outs[0] if len(outs) == 1 else tuple(outs)
