"""OpenVINO Intermediate Representation (`model.xml` + `model.bin`) imported as a torch module.

Reference: `InferenceModel.loadOpenVINO(modelFilePath, weightFilePath, batchSize)` hands the IR to the Inference Engine through JNI
(zoo/.../pipeline/inference/OpenVinoInferenceSupportive.scala:35-57, OpenVINOModel.scala).  There is no Inference Engine here; the
IR itself is a documented container -- an XML graph (`<layers>` with typed layers and ports, `<edges>` between ports) plus one
binary file of constants addressed by (offset, size) -- so it is READ and its layers are evaluated by torch ops, after which the
model runs on the GPU like any other (`InferenceModel.quantize` gives the 8-bit path on FP8 tensor cores).

Two generations of the format are covered:
* IR v10 / v11 (opset1...): constants are `Const` layers, shapes / axes arrive as constant inputs, layer attributes in `<data>`;
* IR v2-v7 (the generation of the reference, OpenVINO 2018-2019): weights and biases hang off the layer as `<blobs>`, layer types
  are the Caffe-like `Convolution`, `FullyConnected`, `Pooling`, `ScaleShift`, `Eltwise`, `Power`, ...
Feed-forward / CNN inference graphs are covered; an unknown layer type raises `NotImplementedError` naming the layer.
"""
from __future__ import annotations

import xml.etree.ElementTree as ET

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

_PREC = {"FP32": np.float32, "F32": np.float32, "FP16": np.float16, "F16": np.float16, "I64": np.int64, "I32": np.int32, "I8": np.int8, "U8": np.uint8,
         "BOOL": np.bool_, "FP64": np.float64, "F64": np.float64}


def _ints(s, default=()):
    return [int(v) for v in s.split(",") if v.strip() != ""] if s not in (None, "") else list(default)


def _dtype(name, default=np.float32):
    return _PREC.get(str(name).upper(), default) if name else default


class _Layer:
    __slots__ = ("id", "name", "type", "version", "data", "in_ports", "out_ports", "blobs", "out_prec")

    def __init__(self, el):
        self.id, self.name, self.type, self.version = el.get("id"), el.get("name"), el.get("type"), el.get("version")
        d = el.find("data")
        self.data = dict(d.attrib) if d is not None else {}
        self.in_ports = [p.get("id") for p in el.findall("input/port")]
        self.out_ports = [p.get("id") for p in el.findall("output/port")]
        self.out_prec = [p.get("precision") or el.get("precision") for p in el.findall("output/port")]
        self.blobs = {b.tag: (int(b.get("offset")), int(b.get("size")), b.get("precision") or el.get("precision")) for b in el.findall("blobs/*")}

    def a(self, key, default=None):
        return self.data.get(key, default)


def _pads(layer, x, kernel, strides, dilations=None):
    """(pads_begin, pads_end) of a spatial op: explicit lists, or computed for auto_pad same_upper / same_lower."""
    n = len(kernel)
    dil = dilations or [1] * n
    auto = (layer.a("auto_pad") or layer.a("auto-pad") or "explicit").lower()
    if auto in ("same_upper", "same_lower"):
        pb, pe = [], []
        for i in range(n):
            size = x.shape[2 + i]
            out = -(-size // strides[i])
            total = max((out - 1) * strides[i] + (kernel[i] - 1) * dil[i] + 1 - size, 0)
            lo, hi = total // 2, total - total // 2
            pb.append(lo if auto == "same_upper" else hi)
            pe.append(hi if auto == "same_upper" else lo)
        return pb, pe
    if auto == "valid":
        return [0] * n, [0] * n
    return _ints(layer.a("pads_begin"), [0] * n), _ints(layer.a("pads_end"), [0] * n)


def _pad_input(x, pb, pe, value=0.0):
    if any(pb) or any(pe):
        flat = []
        for b, e in zip(reversed(pb), reversed(pe)):
            flat += [b, e]
        return F.pad(x, flat, value=value)
    return x


def _conv(layer, x, w, b=None, groups=1):
    n = w.dim() - 2
    strides, dil = _ints(layer.a("strides"), [1] * n), _ints(layer.a("dilations"), [1] * n)
    pb, pe = _pads(layer, x, list(w.shape[2:]), strides, dil)
    fn = {1: F.conv1d, 2: F.conv2d, 3: F.conv3d}[n]
    if pb == pe:
        return fn(x, w, b, strides, pb, dil, groups)
    return fn(_pad_input(x, pb, pe), w, b, strides, 0, dil, groups)


def _pool(layer, x, kind):
    kernel = _ints(layer.a("kernel"))
    n = len(kernel)
    strides = _ints(layer.a("strides"), [1] * n)
    pb, pe = _pads(layer, x, kernel, strides)
    ceil = (layer.a("rounding_type") or layer.a("rounding-type") or "floor").lower() == "ceil"
    if ceil:            # the window may overhang the end: extend the end padding so that floor division gives the ceil size
        for i in range(n):
            size = x.shape[2 + i] + pb[i] + pe[i]
            rem = (size - kernel[i]) % strides[i]
            if rem:
                pe[i] += strides[i] - rem
    if kind == "max":
        fn = {1: F.max_pool1d, 2: F.max_pool2d, 3: F.max_pool3d}[n]
        return fn(_pad_input(x, pb, pe, float("-inf")), kernel, strides)
    exclude = str(layer.a("exclude-pad") or layer.a("exclude_pad") or "false").lower() == "true"
    fn = {1: F.avg_pool1d, 2: F.avg_pool2d, 3: F.avg_pool3d}[n]
    xp = _pad_input(x, pb, pe)
    if not exclude or not (any(pb) or any(pe)):
        return fn(xp, kernel, strides)
    ones = _pad_input(torch.ones_like(x), pb, pe)
    return fn(xp, kernel, strides) / fn(ones, kernel, strides).clamp_min(1e-12)


def _reshape(layer, x, shape):
    shape = [int(v) for v in shape.reshape(-1).tolist()]
    if str(layer.a("special_zero", "false")).lower() == "true":
        shape = [x.shape[i] if v == 0 else v for i, v in enumerate(shape)]
    return x.reshape(shape)


def _fake_quantize(layer, x, in_low, in_high, out_low, out_high):
    levels = int(layer.a("levels", 256))
    scale = (in_high - in_low) / (levels - 1)
    q = torch.round((torch.minimum(torch.maximum(x, in_low), in_high) - in_low) / scale)
    return q * (out_high - out_low) / (levels - 1) + out_low


def _axes(t):
    return [int(v) for v in t.reshape(-1).tolist()]


# ---- IR v10+ ops: fn(layer, *inputs) ---------------------------------------------------------------------------------------------
_OPS10 = {
    "Convolution": lambda l, x, w: _conv(l, x, w),
    "GroupConvolution": lambda l, x, w: _conv(l, x, w.reshape(w.shape[0] * w.shape[1], *w.shape[2:]), None, w.shape[0]),
    "Add": lambda l, a, b: a + b, "Subtract": lambda l, a, b: a - b, "Multiply": lambda l, a, b: a * b, "Divide": lambda l, a, b: a / b,
    "Maximum": lambda l, a, b: torch.maximum(a, b), "Minimum": lambda l, a, b: torch.minimum(a, b), "Power": lambda l, a, b: torch.pow(a, b),
    "Relu": lambda l, x: F.relu(x), "ReLU": lambda l, x: F.relu(x), "Sigmoid": lambda l, x: torch.sigmoid(x), "Tanh": lambda l, x: torch.tanh(x),
    "Exp": lambda l, x: torch.exp(x), "Log": lambda l, x: torch.log(x), "Sqrt": lambda l, x: torch.sqrt(x), "Abs": lambda l, x: torch.abs(x),
    "Negative": lambda l, x: -x, "Erf": lambda l, x: torch.erf(x), "Elu": lambda l, x: F.elu(x, float(l.a("alpha", 1.0))),
    "Gelu": lambda l, x: F.gelu(x), "Swish": lambda l, x, *b: x * torch.sigmoid(x * (b[0] if b else 1.0)), "HSwish": lambda l, x: F.hardswish(x),
    "Mish": lambda l, x: F.mish(x), "HSigmoid": lambda l, x: F.hardsigmoid(x), "SoftPlus": lambda l, x: F.softplus(x),
    "PRelu": lambda l, x, s: torch.where(x >= 0, x, x * (s.reshape(1, -1, *([1] * (x.dim() - 2))) if s.numel() > 1 else s)),
    "Clamp": lambda l, x: torch.clamp(x, float(l.a("min")), float(l.a("max"))),
    "MaxPool": lambda l, x: _pool(l, x, "max"), "AvgPool": lambda l, x: _pool(l, x, "avg"),
    "MatMul": lambda l, a, b: (a.transpose(-1, -2) if str(l.a("transpose_a", "false")).lower() == "true" else a) @
                              (b.transpose(-1, -2) if str(l.a("transpose_b", "false")).lower() == "true" else b),
    "Reshape": _reshape,
    "Squeeze": lambda l, x, *ax: x.squeeze() if not ax else x.reshape([s for i, s in enumerate(x.shape) if i not in {a % x.dim() for a in _axes(ax[0])}]),
    "Unsqueeze": lambda l, x, ax: _unsqueeze(x, _axes(ax)),
    "Transpose": lambda l, x, order: x.permute(_axes(order)) if order.numel() else x.permute(*reversed(range(x.dim()))),
    "Concat": lambda l, *xs: torch.cat(xs, int(l.a("axis", 1))),
    "Softmax": lambda l, x: torch.softmax(x, int(l.a("axis", 1))), "SoftMax": lambda l, x: torch.softmax(x, int(l.a("axis", 1))),
    "LogSoftmax": lambda l, x: torch.log_softmax(x, int(l.a("axis", 1))),
    "ReduceMean": lambda l, x, ax: x.mean(_axes(ax), keepdim=str(l.a("keep_dims", "false")).lower() == "true"),
    "ReduceSum": lambda l, x, ax: x.sum(_axes(ax), keepdim=str(l.a("keep_dims", "false")).lower() == "true"),
    "ReduceMax": lambda l, x, ax: x.amax(_axes(ax), keepdim=str(l.a("keep_dims", "false")).lower() == "true"),
    "Convert": lambda l, x: x.to(torch.from_numpy(np.zeros(0, _dtype(l.a("destination_type")))).dtype),
    "ShapeOf": lambda l, x: torch.tensor(list(x.shape), dtype=torch.int64),
    "Gather": lambda l, x, idx, ax: torch.index_select(x, int(ax), idx.reshape(-1).long()).reshape(
        list(x.shape[:int(ax)]) + list(idx.shape) + list(x.shape[int(ax) + 1:])),
    "FakeQuantize": _fake_quantize,
    "MVN": lambda l, x: _mvn(l, x),
}


def _unsqueeze(x, axes):
    for a in sorted(a if a >= 0 else a + x.dim() + len(axes) for a in axes):
        x = x.unsqueeze(a)
    return x


def _mvn(l, x):
    across = str(l.a("across_channels", "false")).lower() in ("true", "1")
    dims = list(range(1 if across else 2, x.dim()))
    x = x - x.mean(dims, keepdim=True)
    if str(l.a("normalize_variance", "true")).lower() in ("true", "1"):
        x = x / (x.var(dims, unbiased=False, keepdim=True).sqrt() + float(l.a("eps", 1e-9)))
    return x


class OpenVINOModule(nn.Module):
    """The IR graph as a torch module: `forward(*inputs)` in the order of the IR's Parameter / Input layers; one tensor or a tuple
    (in the order of the Result layers, or of the unconsumed outputs for the old format) comes back."""

    def __init__(self, xml_path, bin_path):
        super().__init__()
        root = ET.parse(xml_path).getroot()
        self.version = int(root.get("version", "10"))
        with open(bin_path, "rb") as f:
            self._bin = f.read()
        self.layers = [_Layer(el) for el in root.findall("layers/layer")]
        by_id = {l.id: l for l in self.layers}
        self.src = {}                                       # (to-layer, to-port) -> (from-layer, from-port)
        consumed = set()
        for e in root.findall("edges/edge"):
            self.src[(e.get("to-layer"), e.get("to-port"))] = (e.get("from-layer"), e.get("from-port"))
            consumed.add(e.get("from-layer"))
        self.inputs = [l for l in self.layers if l.type in ("Parameter", "Input")]
        results = [l for l in self.layers if l.type == "Result"]
        self.outputs = results or [l for l in self.layers if l.id not in consumed and l.type not in ("Const",)]
        self._consts = {}
        for k, l in enumerate(self.layers):
            if l.type == "Const":
                arr = self._const_array(l)
                self.register_buffer(f"c{k}", torch.from_numpy(arr))
                self._consts[l.id] = f"c{k}"
            for tag, (off, size, prec) in l.blobs.items():
                if l.type == "Const":
                    continue
                arr = np.frombuffer(self._bin, _dtype(prec), size // np.dtype(_dtype(prec)).itemsize, off).astype(np.float32)
                self.register_buffer(f"b{k}_{tag}", torch.from_numpy(arr.copy()))
        self._index = {l.id: k for k, l in enumerate(self.layers)}
        # matrix products with constant weights become nn.Linear children, so that `InferenceModel.quantize` (FP8) and other module-level
        # tools see them as layers
        self._linear = {}
        for k, l in enumerate(self.layers):
            w = b = None
            if l.type == "MatMul" and self.version >= 10 and len(l.in_ports) == 2 and str(l.a("transpose_a", "false")).lower() != "true":
                s1 = self.src.get((l.id, l.in_ports[1]))
                if s1 is not None and s1[0] in self._consts:
                    c = getattr(self, self._consts[s1[0]])
                    if c.dim() == 2 and c.is_floating_point():
                        w = c if str(l.a("transpose_b", "false")).lower() == "true" else c.t()
            elif l.type == "FullyConnected" and self.version < 10 and "weights" in l.blobs:
                out = int(l.a("out-size"))
                w = getattr(self, f"b{k}_weights").reshape(out, -1)
                b = getattr(self, f"b{k}_biases", None)
            if w is not None:
                lin = nn.Linear(w.shape[1], w.shape[0], bias=b is not None)
                with torch.no_grad():
                    lin.weight.copy_(w)
                    if b is not None:
                        lin.bias.copy_(b)
                name = f"fc_{l.id}"
                self.add_module(name, lin.requires_grad_(False))
                self._linear[l.id] = name
        # evaluation order: layers are already topologically sorted in IR files, but do not rely on it
        order, seen = [], set()

        def visit(lid):
            stack = [(lid, False)]
            while stack:
                cur, done = stack.pop()
                if done:
                    order.append(cur)
                    continue
                if cur in seen:
                    continue
                seen.add(cur)
                stack.append((cur, True))
                for p in by_id[cur].in_ports:
                    s = self.src.get((cur, p))
                    if s is not None and s[0] not in seen:
                        stack.append((s[0], False))

        for o in self.outputs:
            visit(o.id)
        self.order = [by_id[i] for i in order]
        del self._bin

    def _const_array(self, l):
        if l.blobs:                                          # old format: <blobs><custom offset size/></blobs> + output port dims
            off, size, prec = next(iter(l.blobs.values()))
            dt = _dtype(prec)
            arr = np.frombuffer(self._bin, dt, size // np.dtype(dt).itemsize, off)
            return arr.astype(np.float32 if dt == np.float16 else dt).copy()
        dt = _dtype(l.a("element_type"))
        shape = _ints(l.a("shape"))
        off, size = int(l.a("offset", 0)), int(l.a("size", 0))
        arr = np.frombuffer(self._bin, dt, size // np.dtype(dt).itemsize, off).reshape(shape)
        return arr.astype(np.float32 if dt == np.float16 else dt).copy()

    # ---- old-format layers -------------------------------------------------------------------------------------------------------
    def _old(self, l, k, xs):
        blob = lambda tag: getattr(self, f"b{k}_{tag}", None)
        t = l.type
        x = xs[0] if xs else None
        if t == "Convolution":
            kernel = _ints(l.a("kernel")) or [int(l.a("kernel-y")), int(l.a("kernel-x"))]
            out_c, group = int(l.a("output")), int(l.a("group", 1))
            w = blob("weights").reshape(out_c, x.shape[1] // group, *kernel)
            if "strides" not in l.data and "stride-x" in l.data:
                l.data.update(strides=f"{l.a('stride-y', 1)},{l.a('stride-x', 1)}", pads_begin=f"{l.a('pad-y', 0)},{l.a('pad-x', 0)}",
                              pads_end=f"{l.a('pad-b', l.a('pad-y', 0))},{l.a('pad-r', l.a('pad-x', 0))}", dilations=f"{l.a('dilation-y',
                                      1)},{l.a('dilation-x', 1)}")
            return _conv(l, x, w, blob("biases"), group)
        if t == "FullyConnected" and l.id in self._linear:
            return getattr(self, self._linear[l.id])(x.reshape(x.shape[0], -1))
        if t == "FullyConnected":
            out = int(l.a("out-size"))
            x2 = x.reshape(x.shape[0], -1)
            return F.linear(x2, blob("weights").reshape(out, x2.shape[1]), blob("biases"))
        if t == "Pooling":
            if "kernel" not in l.data:
                l.data.update(kernel=f"{l.a('kernel-y')},{l.a('kernel-x')}", strides=f"{l.a('stride-y', 1)},{l.a('stride-x', 1)}",
                              pads_begin=f"{l.a('pad-y', 0)},{l.a('pad-x', 0)}", pads_end=f"{l.a('pad-b', l.a('pad-y', 0))},{l.a('pad-r', l.a('pad-x', 0))}")
            return _pool(l, x, "max" if (l.a("pool-method") or "max").lower() == "max" else "avg")
        if t == "ScaleShift":
            shape = [1, -1] + [1] * (x.dim() - 2)
            y = x * blob("weights").reshape(shape)
            return y + blob("biases").reshape(shape) if blob("biases") is not None else y
        if t == "Eltwise":
            op = (l.a("operation") or "sum").lower()
            y = xs[0]
            for o in xs[1:]:
                y = {"sum": torch.add, "mul": torch.mul, "max": torch.maximum, "sub": torch.sub, "min": torch.minimum, "div": torch.div}[op](y, o)
            return y
        if t == "Power":
            return torch.pow(x * float(l.a("scale", 1.0)) + float(l.a("shift", 0.0)), float(l.a("power", 1.0)))
        if t == "ReLU":
            return F.leaky_relu(x, float(l.a("negative_slope", 0.0))) if float(l.a("negative_slope", 0.0)) else F.relu(x)
        if t == "Activation":
            return {"relu": F.relu, "sigmoid": torch.sigmoid, "tanh": torch.tanh, "elu": F.elu, "exp": torch.exp}[(l.a("type") or "relu").lower()](x)
        if t == "Norm":                                       # LRN
            size = int(l.a("local-size"))
            if (l.a("region") or "across").lower() != "across":
                raise NotImplementedError("Norm within a channel")
            return F.local_response_norm(x, size, float(l.a("alpha")), float(l.a("beta")), float(l.a("k", 1.0)))
        if t == "Reshape" and len(xs) == 1:
            return x.reshape([x.shape[i] if v == 0 else v for i, v in enumerate(_ints(l.a("dim")))])
        if t == "Flatten":
            return x.flatten(int(l.a("axis", 1)))
        if t == "Permute":
            return x.permute(_ints(l.a("order")))
        if t == "BatchNormalization":
            return F.batch_norm(x, blob("mean") if blob("mean") is not None else blob("weights"),
                    blob("variance") if blob("variance") is not None else blob("biases"),
                                None, None, False, 0.0, float(l.a("epsilon", 1e-5)))
        fn = _OPS10.get(t)
        if fn is None:
            raise NotImplementedError(f"OpenVINO IR layer type {t!r} (layer {l.name!r}) is not supported by the importer")
        return fn(l, *xs)

    def forward(self, *xs):
        if len(xs) == 1 and isinstance(xs[0], (tuple, list)):
            xs = tuple(xs[0])
        if len(xs) != len(self.inputs):
            raise ValueError(f"expected {len(self.inputs)} inputs ({[l.name for l in self.inputs]}), got {len(xs)}")
        vals = {}
        for l, x in zip(self.inputs, xs):
            vals[(l.id, l.out_ports[0])] = x if torch.is_tensor(x) else torch.as_tensor(np.asarray(x))
        dev = xs[0].device if torch.is_tensor(xs[0]) else None
        outs = []
        for l in self.order:
            if l.type in ("Parameter", "Input"):
                continue
            if l.type == "Const":
                vals[(l.id, l.out_ports[0])] = getattr(self, self._consts[l.id])
                continue
            args = [vals[self.src[(l.id, p)]] for p in l.in_ports]
            if l.type == "Result":
                vals[(l.id, "r")] = args[0]
                continue
            if dev is not None:
                args = [a.to(dev) if torch.is_tensor(a) and a.is_floating_point() and a.device != dev else a for a in args]
            y = self._old(l, self._index[l.id], args) if self.version < 10 else self._new(l, args)
            ys = y if isinstance(y, tuple) else (y,)
            for p, v in zip(l.out_ports, ys):
                vals[(l.id, p)] = v
        for o in self.outputs:
            outs.append(vals[(o.id, "r")] if o.type == "Result" else vals[(o.id, o.out_ports[0])])
        return outs[0] if len(outs) == 1 else tuple(outs)

    def _new(self, l, args):
        if l.id in self._linear:
            return getattr(self, self._linear[l.id])(args[0])
        fn = _OPS10.get(l.type)
        if fn is None:
            raise NotImplementedError(f"OpenVINO IR op {l.type!r} (layer {l.name!r}, {l.version}) is not supported by the importer")
        return fn(l, *args)


def load_openvino_ir(xml_path, bin_path=None):
    """`model.xml` (+ `model.bin`, default: same name) -> OpenVINOModule in eval mode."""
    import os

    if bin_path is None:
        bin_path = os.path.splitext(xml_path)[0] + ".bin"
    return OpenVINOModule(xml_path, bin_path).eval()


def save_openvino_ir(model, xml_path, input_shape, bin_path=None, name="model"):
    """Write a feed-forward torch model (Sequential of Conv2d / Linear / ReLU / Sigmoid / Tanh / MaxPool2d / AvgPool2d / Flatten /
    Softmax / Dropout) as IR v10 (`xml_path` + `.bin`): what OpenVINO's Model Optimizer would emit for it, so that a model trained here
    can be handed to an OpenVINO deployment, and the form `load_openvino_ir` reads back."""
    import os

    bin_path = bin_path or os.path.splitext(xml_path)[0] + ".bin"
    blob, layers, edges = bytearray(), [], []
    counter = {"id": 0}

    def port(i, shape=()):
        return f'<port id="{i}" precision="FP32">' + "".join(f"<dim>{int(d)}</dim>" for d in shape) + "</port>"

    def add(type_, lname, data="", n_in=0, shape=()):
        i = counter["id"]
        counter["id"] += 1
        ins = "<input>" + "".join(port(k) for k in range(n_in)) + "</input>" if n_in else ""
        outs = "" if type_ == "Result" else f"<output>{port(n_in, shape)}</output>"
        layers.append(f'<layer id="{i}" name="{lname}" type="{type_}" version="opset1">{data}{ins}{outs}</layer>')
        return i, n_in

    def const(arr, lname, dtype=np.float32):
        a = np.ascontiguousarray(np.asarray(arr, dtype=dtype))
        off = len(blob)
        blob.extend(a.tobytes())
        et = {np.dtype(np.float32): "f32", np.dtype(np.int64): "i64"}[a.dtype]
        return add("Const", lname, f'<data element_type="{et}" offset="{off}" shape="{",".join(map(str, a.shape))}" size="{a.nbytes}"/>', 0, a.shape)

    def connect(src, dst, to_port):
        edges.append(f'<edge from-layer="{src[0]}" from-port="{src[1]}" to-layer="{dst[0]}" to-port="{to_port}"/>')

    def binary(type_, lname, a, b, data=""):
        node = add(type_, lname, data, 2)
        connect(a, node, 0)
        connect(b, node, 1)
        return node

    def unary(type_, lname, a, data=""):
        node = add(type_, lname, data, 1)
        connect(a, node, 0)
        return node

    mods = list(model) if isinstance(model, nn.Sequential) else [model]
    cur = add("Parameter", "data", f'<data element_type="f32" shape="{",".join(map(str, input_shape))}"/>', 0, input_shape)
    pair = lambda v: (v, v) if isinstance(v, int) else tuple(v)
    for k, m in enumerate(mods):
        n = f"{type(m).__name__.lower()}_{k}"
        if isinstance(m, nn.Conv2d):
            w = m.weight.detach().float().cpu().numpy()
            attrs = (f'<data auto_pad="explicit" dilations="{m.dilation[0]},{m.dilation[1]}" pads_begin="{m.padding[0]},{m.padding[1]}" '
                     f'pads_end="{m.padding[0]},{m.padding[1]}" strides="{m.stride[0]},{m.stride[1]}"/>')
            if m.groups == 1:
                cur = binary("Convolution", n, cur, const(w, n + "_w"), attrs)
            else:
                cur = binary("GroupConvolution", n, cur, const(w.reshape(m.groups, w.shape[0] // m.groups, *w.shape[1:]), n + "_w"), attrs)
            if m.bias is not None:
                cur = binary("Add", n + "_bias", cur, const(m.bias.detach().float().cpu().numpy().reshape(1, -1, 1, 1), n + "_b"))
        elif isinstance(m, nn.Linear):
            cur = binary("MatMul", n, cur, const(m.weight.detach().float().cpu().numpy(), n + "_w"), '<data transpose_a="false" transpose_b="true"/>')
            if m.bias is not None:
                cur = binary("Add", n + "_bias", cur, const(m.bias.detach().float().cpu().numpy().reshape(1, -1), n + "_b"))
        elif isinstance(m, (nn.ReLU, nn.Sigmoid, nn.Tanh)):
            cur = unary({"ReLU": "Relu", "Sigmoid": "Sigmoid", "Tanh": "Tanh"}[type(m).__name__], n, cur)
        elif isinstance(m, (nn.MaxPool2d, nn.AvgPool2d)):
            ks, st, pd = pair(m.kernel_size), pair(m.stride or m.kernel_size), pair(m.padding)
            extra = "" if isinstance(m, nn.MaxPool2d) else f' exclude-pad="{str(not m.count_include_pad).lower()}"'
            cur = unary("MaxPool" if isinstance(m, nn.MaxPool2d) else "AvgPool", n, cur,
                        f'<data auto_pad="explicit" kernel="{ks[0]},{ks[1]}" pads_begin="{pd[0]},{pd[1]}" pads_end="{pd[0]},{pd[1]}" '
                        f'rounding_type="{"ceil" if m.ceil_mode else "floor"}" strides="{st[0]},{st[1]}"{extra}/>')
        elif isinstance(m, nn.Flatten):
            cur = binary("Reshape", n, cur, const([0, -1], n + "_shape", np.int64), '<data special_zero="true"/>')
        elif isinstance(m, nn.Softmax):
            cur = unary("Softmax", n, cur, f'<data axis="{1 if m.dim in (-1, None) else m.dim}"/>')
        elif isinstance(m, (nn.Dropout, nn.Identity)):
            continue
        else:
            raise NotImplementedError(f"save_openvino_ir: no IR op written for layer {type(m).__name__}")
    connect(cur, add("Result", "result", "", 1), 0)
    with open(xml_path, "w") as f:
        f.write(f'<?xml version="1.0" ?>\n<net name="{name}" '
                f'version="10">\n<layers>\n' + "\n".join(layers) + "\n</layers>\n<edges>\n" + "\n".join(edges) + "\n</edges>\n</net>\n")
    with open(bin_path, "wb") as f:
        f.write(bytes(blob))
    return xml_path, bin_path
