"""ONNX loader (reference: pyzoo/zoo/pipeline/api/onnx/onnx_loader.py + 42 operator mappers under mapper/).

`OnnxLoader.from_path(path)` / `OnnxLoader(graph_proto).to_keras()` turn an ONNX graph into a trainable model.  The
reference maps every node to a zoo keras layer; here the graph is kept as a graph: `OnnxModel` is a `KerasNet` that
executes the nodes in order with torch ops, initializers become parameters (so the model can be fine-tuned), and
the same 42 operators are covered:
Abs Add AveragePool BatchNormalization Cast Clip Concat Constant Conv Div Dropout Elu Exp Flatten Gather Gemm
GlobalAveragePool Greater HardSigmoid LeakyRelu Log LogSoftmax LRN MatMul MaxPool Mul Neg Pow ReduceMean ReduceSum
Relu Reshape Shape Sigmoid Slice Softmax Sqrt Squeeze Sub Tanh Transpose Unsqueeze  (+ Identity).
Reading a `.onnx` file needs the `onnx` package for its protobuf schema; it is imported lazily and its absence is
reported as such.  Any object with the ModelProto / GraphProto attribute layout is accepted.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ..keras.engine.topology import KerasNet

_ONNX_DTYPES = {1: np.float32, 2: np.uint8, 3: np.int8, 5: np.int16, 6: np.int32, 7: np.int64, 9: np.bool_, 10: np.float16, 11: np.float64}


def _tensor_to_numpy(t):
    if isinstance(t, np.ndarray):
        return t
    if hasattr(t, "numpy_value"):
        return np.asarray(t.numpy_value)
    try:
        from onnx import numpy_helper

        return numpy_helper.to_array(t)
    except ImportError:
        pass
    dt = _ONNX_DTYPES.get(getattr(t, "data_type", 1), np.float32)
    dims = tuple(getattr(t, "dims", ()))
    raw = getattr(t, "raw_data", b"")
    if raw:
        return np.frombuffer(raw, dtype=dt).reshape(dims).copy()
    for field in ("float_data", "int64_data", "int32_data", "double_data"):
        v = list(getattr(t, field, []) or [])
        if v:
            return np.asarray(v, dtype=dt).reshape(dims)
    return np.zeros(dims, dtype=dt)


def _attrs(node):
    out = {}
    for a in getattr(node, "attribute", []) or []:
        if isinstance(a, tuple):
            out[a[0]] = a[1]
            continue
        t = getattr(a, "type", None)
        if t == 1 or (t is None and hasattr(a, "f") and not getattr(a, "ints", None)):
            out[a.name] = a.f
        elif t == 2:
            out[a.name] = a.i
        elif t == 3:
            out[a.name] = a.s.decode() if isinstance(a.s, bytes) else a.s
        elif t == 4:
            out[a.name] = _tensor_to_numpy(a.t)
        elif t == 6:
            out[a.name] = list(a.floats)
        elif t == 7:
            out[a.name] = list(a.ints)
        else:
            out[a.name] = getattr(a, "value", None)
    return out


def _pool_pads(pads, nd):
    if not pads:
        return [0] * nd, None
    begin, end = list(pads[:nd]), list(pads[nd:])
    if begin == end:
        return begin, None
    extra = []
    for b, e in zip(reversed(begin), reversed(end)):
        extra += [b, e]
    return [0] * nd, extra


class OnnxModel(KerasNet):
    def __init__(self, graph):
        super().__init__()
        self.graph = graph
        inits = {i.name: _tensor_to_numpy(i) for i in graph.initializer}
        self.input_names = [i.name for i in graph.input if i.name not in inits]
        self.output_names = [o.name for o in graph.output]
        self._pnames = {}
        self.params = nn.ParameterDict()
        self.consts = {}
        for k, (name, arr) in enumerate(inits.items()):
            t = torch.from_numpy(np.ascontiguousarray(arr))
            if t.is_floating_point():
                key = f"p{k}"
                self.params[key] = nn.Parameter(t.float())
                self._pnames[name] = key
            else:
                self.consts[name] = t
        self.nodes = [(n.op_type, list(n.input), list(n.output), _attrs(n)) for n in graph.node]
        unsupported = sorted({op for op, *_ in self.nodes if not hasattr(self, "_op_" + op)})
        if unsupported:
            raise NotImplementedError(f"ONNX operators without a mapper: {unsupported}")

    def forward(self, *xs):
        if len(xs) != len(self.input_names):
            raise ValueError(f"model expects {len(self.input_names)} inputs {self.input_names}, got {len(xs)}")
        env = dict(zip(self.input_names, xs))
        dev = xs[0].device
        for name, key in self._pnames.items():
            env[name] = self.params[key]
        for name, t in self.consts.items():
            env[name] = t.to(dev)
        for op, ins, outs, at in self.nodes:
            args = [env[i] if i else None for i in ins]
            res = getattr(self, "_op_" + op)(args, at)
            res = res if isinstance(res, (tuple, list)) else [res]
            for o, r in zip(outs, res):
                env[o] = r
        ys = [env[o] for o in self.output_names]
        return ys[0] if len(ys) == 1 else tuple(ys)

    # ---- operators -------------------------------------------------------------------------------------
    @staticmethod
    def _i(t):
        return [int(v) for v in (t.tolist() if torch.is_tensor(t) else t)]

    def _op_Abs(self, a, at): return a[0].abs()
    def _op_Add(self, a, at): return a[0] + a[1]
    def _op_Sub(self, a, at): return a[0] - a[1]
    def _op_Mul(self, a, at): return a[0] * a[1]
    def _op_Div(self, a, at): return a[0] / a[1]
    def _op_Neg(self, a, at): return -a[0]
    def _op_Exp(self, a, at): return a[0].exp()
    def _op_Log(self, a, at): return a[0].log()
    def _op_Sqrt(self, a, at): return a[0].sqrt()
    def _op_Pow(self, a, at): return a[0].pow(a[1])
    def _op_Relu(self, a, at): return F.relu(a[0])
    def _op_Sigmoid(self, a, at): return torch.sigmoid(a[0])
    def _op_Tanh(self, a, at): return torch.tanh(a[0])
    def _op_Identity(self, a, at): return a[0]
    def _op_Elu(self, a, at): return F.elu(a[0], at.get("alpha", 1.0))
    def _op_LeakyRelu(self, a, at): return F.leaky_relu(a[0], at.get("alpha", 0.01))
    def _op_HardSigmoid(self, a, at): return torch.clamp(at.get("alpha", 0.2) * a[0] + at.get("beta", 0.5), 0, 1)
    def _op_Softmax(self, a, at): return torch.softmax(a[0], at.get("axis", -1))
    def _op_LogSoftmax(self, a, at): return torch.log_softmax(a[0], at.get("axis", -1))
    def _op_Greater(self, a, at): return a[0] > a[1]
    def _op_Dropout(self, a, at): return F.dropout(a[0], at.get("ratio", 0.5), self.training)
    def _op_Flatten(self, a, at): return a[0].flatten(at.get("axis", 1)) if at.get("axis", 1) else a[0].reshape(1, -1)
    def _op_MatMul(self, a, at): return torch.matmul(a[0], a[1])
    def _op_Concat(self, a, at): return torch.cat(a, at.get("axis", 1))
    def _op_Constant(self, a, at): return torch.from_numpy(np.asarray(at["value"]))
    def _op_Shape(self, a, at): return torch.tensor(list(a[0].shape), dtype=torch.int64)
    def _op_Transpose(self, a, at): return a[0].permute(*at["perm"]) if "perm" in at else a[0].permute(*reversed(range(a[0].dim())))
    def _op_GlobalAveragePool(self, a, at): return a[0].mean(tuple(range(2, a[0].dim())), keepdim=True)
    def _op_LRN(self, a, at): return F.local_response_norm(a[0], at["size"], at.get("alpha", 1e-4), at.get("beta", 0.75), at.get("bias", 1.0))

    def _op_Cast(self, a, at):
        return a[0].to({1: torch.float32, 6: torch.int32, 7: torch.int64, 9: torch.bool, 10: torch.float16, 11: torch.float64}[at["to"]])

    def _op_Clip(self, a, at):
        lo = at.get("min", a[1] if len(a) > 1 and a[1] is not None else None)
        hi = at.get("max", a[2] if len(a) > 2 and a[2] is not None else None)
        return torch.clamp(a[0], None if lo is None else float(lo), None if hi is None else float(hi))

    def _op_Gemm(self, a, at):
        x, w = a[0], a[1]
        x = x.t() if at.get("transA", 0) else x
        w = w.t() if at.get("transB", 0) else w
        y = at.get("alpha", 1.0) * (x @ w)
        return y + at.get("beta", 1.0) * a[2] if len(a) > 2 and a[2] is not None else y

    def _op_Conv(self, a, at):
        x, w, b = a[0], a[1], (a[2] if len(a) > 2 else None)
        nd = w.dim() - 2
        pads, extra = _pool_pads(at.get("pads"), nd)
        if extra:
            x = F.pad(x, extra)
        fn = {1: F.conv1d, 2: F.conv2d, 3: F.conv3d}[nd]
        return fn(x, w, b, at.get("strides", [1] * nd), pads, at.get("dilations", [1] * nd), at.get("group", 1))

    def _pool(self, a, at, kind):
        x = a[0]
        k = at["kernel_shape"]
        nd = len(k)
        pads, extra = _pool_pads(at.get("pads"), nd)
        if extra:
            x = F.pad(x, extra, value=float("-inf") if kind == "max" else 0.0)
        fn = {("max", 1): F.max_pool1d, ("max", 2): F.max_pool2d, ("max", 3): F.max_pool3d,
              ("avg", 1): F.avg_pool1d, ("avg", 2): F.avg_pool2d, ("avg", 3): F.avg_pool3d}[(kind, nd)]
        kw = dict(ceil_mode=bool(at.get("ceil_mode", 0)))
        if kind == "avg":
            kw["count_include_pad"] = bool(at.get("count_include_pad", 0))
        return fn(x, k, at.get("strides", [1] * nd), pads, **kw)

    def _op_MaxPool(self, a, at): return self._pool(a, at, "max")
    def _op_AveragePool(self, a, at): return self._pool(a, at, "avg")

    def _op_BatchNormalization(self, a, at):
        x, scale, bias, mean, var = a[:5]
        return F.batch_norm(x, mean.detach(), var.detach(), scale, bias, False, 0.0, at.get("epsilon", 1e-5))

    def _op_Reshape(self, a, at):
        shape = at.get("shape") or self._i(a[1])
        shape = [a[0].shape[i] if s == 0 else s for i, s in enumerate(shape)]
        return a[0].reshape(shape)

    def _op_Squeeze(self, a, at):
        axes = at.get("axes") or (self._i(a[1]) if len(a) > 1 and a[1] is not None else None)
        x = a[0]
        if axes is None:
            return x.squeeze()
        for ax in sorted([ax % x.dim() for ax in axes], reverse=True):
            x = x.squeeze(ax)
        return x

    def _op_Unsqueeze(self, a, at):
        axes = at.get("axes") or self._i(a[1])
        x = a[0]
        for ax in sorted(axes):
            x = x.unsqueeze(ax)
        return x

    def _op_Gather(self, a, at):
        idx = a[1].long()
        return torch.index_select(a[0], at.get("axis", 0), idx.reshape(-1)).reshape(
            *a[0].shape[:at.get("axis", 0)], *idx.shape, *a[0].shape[at.get("axis", 0) + 1:])

    def _op_Slice(self, a, at):
        x = a[0]
        if "starts" in at:
            starts, ends, axes, steps = at["starts"], at["ends"], at.get("axes", list(range(len(at["starts"])))), [1] * len(at["starts"])
        else:
            starts, ends = self._i(a[1]), self._i(a[2])
            axes = self._i(a[3]) if len(a) > 3 and a[3] is not None else list(range(len(starts)))
            steps = self._i(a[4]) if len(a) > 4 and a[4] is not None else [1] * len(starts)
        sl = [slice(None)] * x.dim()
        for s, e, ax, st in zip(starts, ends, axes, steps):
            n = x.shape[ax]
            sl[ax] = slice(max(s + n, 0) if s < 0 else min(s, n), max(e + n, 0) if e < 0 else min(e, n), st)
        return x[tuple(sl)]

    def _reduce(self, a, at, fn):
        axes = at.get("axes") or (self._i(a[1]) if len(a) > 1 and a[1] is not None else list(range(a[0].dim())))
        return fn(a[0], tuple(axes), keepdim=bool(at.get("keepdims", 1)))

    def _op_ReduceMean(self, a, at): return self._reduce(a, at, torch.mean)
    def _op_ReduceSum(self, a, at): return self._reduce(a, at, torch.sum)


class OnnxLoader:
    def __init__(self, onnx_graph):
        self.graph = onnx_graph.graph if hasattr(onnx_graph, "graph") else onnx_graph

    @classmethod
    def from_path(cls, onnx_path, is_training=False):
        try:
            import onnx
        except ImportError as e:
            raise ImportError("reading .onnx files needs the `onnx` package (its protobuf schema), which is not installed on this box; "
                              "pass an already parsed ModelProto to OnnxLoader(...) instead") from e
        return cls(onnx.load(onnx_path)).to_keras()

    @staticmethod
    def run_node(node, inputs, is_training=False):
        """Execute a single NodeProto on ndarrays (the reference's per-mapper unit-test entry point)."""
        m = OnnxModel.__new__(OnnxModel)
        nn.Module.__init__(m)
        m.training = is_training
        res = getattr(m, "_op_" + node.op_type)([torch.as_tensor(np.asarray(i)) for i in inputs], _attrs(node))
        return [r.detach().numpy() for r in (res if isinstance(res, (tuple, list)) else [res])]

    def to_keras(self):
        return OnnxModel(self.graph)


class OnnxInput:
    def __init__(self, name, zvalue, data=None):
        self.name, self.zvalue, self.data = name, zvalue, data
