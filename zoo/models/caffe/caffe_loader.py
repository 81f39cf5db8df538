"""Caffe model import: `.prototxt` (+ optional `.caffemodel`) -> a torch module graph.

Reference: zoo/src/main/scala/com/intel/analytics/zoo/models/caffe/{CaffeLoader,Converter,LayerConverter,V1LayerConverter}.scala
(2,898 lines over generated protobuf classes) and `Net.loadCaffe` (pipeline/api/Net.scala).  Same behaviour:

* the network definition comes from the prototxt, the weights from the binary `NetParameter` of the caffemodel, matched BY LAYER
  NAME (`matchAll`: a layer that has parameters but no weights in the file is an error);
* V2 (`layer { type: "Convolution" }`) and V1 (`layers { type: CONVOLUTION }`) definitions are both understood;
* data / loss / accuracy layers are dropped, in-place layers (top == bottom) are chained, blobs nobody consumes are the outputs;
* `customized_converters` maps an upper-case layer type to a function `(layer_dict, blobs) -> nn.Module` for user layers;
* `CaffeLoader.load(model, def, weights)` copies the weights into an EXISTING torch model by module name.

The files are read with :mod:`zoo.util.protowire` (no caffe / protobuf-generated classes needed).  The converted network is an
ordinary `nn.Module`, so it trains on this framework's engine and serves through `InferenceModel`.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ...util import protowire as pw
from . import caffe_schema as cs


class CaffeConversionException(Exception):
    pass


def _one(d, k, default=None):
    v = d.get(k)
    return v[-1] if v else default


def _sub(d, k):
    v = d.get(k)
    return v[-1] if v else {}


def _blob_array(b):
    data = pw.floats(b, "data")
    if data.size == 0 and b.get("double_data"):
        data = np.concatenate([np.atleast_1d(np.asarray(p, dtype=np.float64)) for p in b["double_data"]]).astype(np.float32)
    shape = _sub(b, "shape").get("dim")
    if not shape:
        shape = [int(_one(b, k, 1)) for k in ("num", "channels", "height", "width")]
        while len(shape) > 1 and shape[0] == 1 and int(np.prod(shape[1:])) == data.size:      # legacy 4-D blobs of vectors / matrices
            shape = shape[1:]
    return torch.from_numpy(np.array(data, dtype=np.float32).reshape([int(s) for s in shape]))


# ------------------------------------------------------------------------------------------------------ layer modules --
class _Lambda(nn.Module):
    def __init__(self, fn, desc=""):
        super().__init__()
        self.fn, self.desc = fn, desc

    def forward(self, *xs):
        return self.fn(*xs)

    def extra_repr(self):
        return self.desc


class _CaffePool(nn.Module):
    """Caffe rounds pooled sizes UP (ceil mode) and clips the last window so that it starts inside the padded input."""

    def __init__(self, kind, k, s, p, global_pooling):
        super().__init__()
        self.kind, self.k, self.s, self.p, self.global_pooling = kind, k, s, p, global_pooling

    def forward(self, x):
        if self.global_pooling:
            return F.adaptive_max_pool2d(x, 1) if self.kind == 0 else F.adaptive_avg_pool2d(x, 1)
        if self.kind == 0:
            return F.max_pool2d(x, self.k, self.s, self.p, ceil_mode=True)
        return F.avg_pool2d(x, self.k, self.s, self.p, ceil_mode=True, count_include_pad=True)


class _Scale(nn.Module):
    def __init__(self, channels, bias):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(channels))
        self.bias = nn.Parameter(torch.zeros(channels)) if bias else None

    def forward(self, x):
        shape = [1, -1] + [1] * (x.dim() - 2)
        y = x * self.weight.view(shape)
        return y + self.bias.view(shape) if self.bias is not None else y


class _InnerProduct(nn.Linear):
    """Caffe flattens everything from `axis` on before the product."""

    def forward(self, x):
        return super().forward(x.flatten(1))


def _hw(p, base, h, w, default):
    v = p.get(base)
    if v:
        v = [int(x) for x in v]
        return (v[0], v[-1]) if len(v) > 1 else (v[0], v[0])
    if h in p or w in p:
        return int(_one(p, h, default)), int(_one(p, w, default))
    return default, default


def _convert(ltype, layer, blobs, in_channels):
    """Returns (module, out_channels or None).  `blobs` are the layer's weight tensors (may be empty: definition only)."""
    T = ltype.upper().replace("_", "")
    if T in ("CONVOLUTION", "DECONVOLUTION"):
        p = _sub(layer, "convolution_param")
        n_out, group = int(_one(p, "num_output")), int(_one(p, "group", 1))
        k = _hw(p, "kernel_size", "kernel_h", "kernel_w", 1)
        s = _hw(p, "stride", "stride_h", "stride_w", 1)
        pad = _hw(p, "pad", "pad_h", "pad_w", 0)
        dil = _hw(p, "dilation", "_", "_", 1)
        bias = bool(_one(p, "bias_term", True))
        cin = in_channels
        if blobs:
            cin = blobs[0].shape[1] * group if T == "CONVOLUTION" else blobs[0].shape[0]
        if cin is None:
            raise CaffeConversionException(f"layer {_one(layer, 'name')}: cannot infer the input channels without weights or an input shape")
        if T == "CONVOLUTION":
            m = nn.Conv2d(cin, n_out, k, s, pad, dil, group, bias)
        else:
            m = nn.ConvTranspose2d(cin, n_out, k, s, pad, groups=group, bias=bias, dilation=dil)
        if blobs:
            with torch.no_grad():
                m.weight.copy_(blobs[0].reshape(m.weight.shape))
                if bias and len(blobs) > 1:
                    m.bias.copy_(blobs[1].reshape(-1))
        return m, n_out
    if T == "INNERPRODUCT":
        p = _sub(layer, "inner_product_param")
        n_out, bias = int(_one(p, "num_output")), bool(_one(p, "bias_term", True))
        if not blobs:
            raise CaffeConversionException(f"layer {_one(layer, 'name')}: InnerProduct needs its weights to know the input width")
        w = blobs[0].reshape(n_out, -1) if not _one(p, "transpose", False) else blobs[0].reshape(-1, n_out).t()
        m = _InnerProduct(w.shape[1], n_out, bias)
        with torch.no_grad():
            m.weight.copy_(w)
            if bias and len(blobs) > 1:
                m.bias.copy_(blobs[1].reshape(-1))
        return m, n_out
    if T == "RELU":
        slope = float(_one(_sub(layer, "relu_param"), "negative_slope", 0.0))
        return (nn.LeakyReLU(slope) if slope else nn.ReLU()), in_channels
    if T == "PRELU":
        shared = bool(_one(_sub(layer, "prelu_param"), "channel_shared", False))
        m = nn.PReLU(1 if shared or not blobs else blobs[0].numel())
        if blobs:
            with torch.no_grad():
                m.weight.copy_(blobs[0].reshape(-1))
        return m, in_channels
    if T == "ELU":
        return nn.ELU(float(_one(_sub(layer, "elu_param"), "alpha", 1.0))), in_channels
    if T in ("TANH", "SIGMOID", "ABSVAL", "BNLL"):
        return {"TANH": nn.Tanh(), "SIGMOID": nn.Sigmoid(), "ABSVAL": _Lambda(torch.abs, "abs"), "BNLL": nn.Softplus()}[T], in_channels
    if T == "POOLING":
        p = _sub(layer, "pooling_param")
        kind = _one(p, "pool", 0)
        kind = {"MAX": 0, "AVE": 1, "STOCHASTIC": 2}.get(kind, kind) if isinstance(kind, str) else int(kind)
        if kind == 2:
            raise CaffeConversionException("stochastic pooling is not supported")
        k = _hw(p, "kernel_size", "kernel_h", "kernel_w", 1)
        s = _hw(p, "stride", "stride_h", "stride_w", 1)
        pad = _hw(p, "pad", "pad_h", "pad_w", 0)
        return _CaffePool(kind, k, s, pad, bool(_one(p, "global_pooling", False))), in_channels
    if T == "LRN":
        p = _sub(layer, "lrn_param")
        region = _one(p, "norm_region", 0)
        if region in (1, "WITHIN_CHANNEL"):
            raise CaffeConversionException("LRN within channel is not supported")
        return nn.LocalResponseNorm(int(_one(p, "local_size", 5)), float(_one(p, "alpha", 1.0)), float(_one(p, "beta", 0.75)),
                                    float(_one(p, "k", 1.0))), in_channels
    if T == "DROPOUT":
        return nn.Dropout(float(_one(_sub(layer, "dropout_param"), "dropout_ratio", 0.5))), in_channels
    if T in ("SOFTMAX", "SOFTMAXLOSS", "SOFTMAXWITHLOSS"):
        return nn.Softmax(int(_one(_sub(layer, "softmax_param"), "axis", 1))), in_channels
    if T == "SIGMOIDCROSSENTROPYLOSS":
        return nn.Sigmoid(), in_channels
    if T == "BATCHNORM":
        p = _sub(layer, "batch_norm_param")
        ch = blobs[0].numel() if blobs else in_channels
        m = nn.BatchNorm2d(ch, eps=float(_one(p, "eps", 1e-5)), momentum=1.0 - float(_one(p, "moving_average_fraction", 0.999)), affine=False)
        if blobs:
            f = float(blobs[2].reshape(-1)[0]) if len(blobs) > 2 else 1.0
            f = 1.0 / f if f != 0 else 0.0              # caffe stores un-normalised sums and a scale factor
            with torch.no_grad():
                m.running_mean.copy_(blobs[0].reshape(-1) * f)
                m.running_var.copy_(blobs[1].reshape(-1) * f)
        return m, ch
    if T in ("SCALE", "BIAS"):
        ch = blobs[0].numel() if blobs else in_channels
        has_bias = T == "BIAS" or bool(_one(_sub(layer, "scale_param"), "bias_term", False))
        m = _Scale(ch, has_bias)
        if blobs:
            with torch.no_grad():
                if T == "SCALE":
                    m.weight.copy_(blobs[0].reshape(-1))
                    if has_bias and len(blobs) > 1:
                        m.bias.copy_(blobs[1].reshape(-1))
                else:
                    m.bias.copy_(blobs[0].reshape(-1))
        return m, ch
    if T == "CONCAT":
        p = _sub(layer, "concat_param")
        axis = int(_one(p, "axis", _one(p, "concat_dim", 1)))
        return _Lambda(lambda *xs: torch.cat(xs, axis), f"concat(axis={axis})"), None
    if T == "ELTWISE":
        p = _sub(layer, "eltwise_param")
        op = _one(p, "operation", 1)
        op = {"PROD": 0, "SUM": 1, "MAX": 2}.get(op, op) if isinstance(op, str) else int(op)
        coeff = [float(c) for c in pw.floats(p, "coeff")] if p.get("coeff") else None

        def elt(*xs):
            if op == 0:
                y = xs[0]
                for x in xs[1:]:
                    y = y * x
                return y
            if op == 2:
                y = xs[0]
                for x in xs[1:]:
                    y = torch.maximum(y, x)
                return y
            cs_ = coeff or [1.0] * len(xs)
            y = xs[0] * cs_[0]
            for c, x in zip(cs_[1:], xs[1:]):
                y = y + c * x
            return y

        return _Lambda(elt, f"eltwise(op={op})"), in_channels
    if T == "FLATTEN":
        p = _sub(layer, "flatten_param")
        return nn.Flatten(int(_one(p, "axis", 1)), int(_one(p, "end_axis", -1))), None
    if T == "RESHAPE":
        dims = [int(d) for d in _sub(_sub(layer, "reshape_param"), "shape").get("dim", [])]

        def reshape(x):
            shape = [x.shape[i] if d == 0 else d for i, d in enumerate(dims)]
            return x.reshape(shape)

        return _Lambda(reshape, f"reshape({dims})"), None
    if T == "POWER":
        p = _sub(layer, "power_param")
        pw_, sc, sh = float(_one(p, "power", 1.0)), float(_one(p, "scale", 1.0)), float(_one(p, "shift", 0.0))
        return _Lambda(lambda x: torch.pow(sh + sc * x, pw_), f"power({pw_},{sc},{sh})"), in_channels
    if T in ("EXP", "LOG"):
        p = _sub(layer, "exp_param" if T == "EXP" else "log_param")
        base, sc, sh = float(_one(p, "base", -1.0)), float(_one(p, "scale", 1.0)), float(_one(p, "shift", 0.0))
        lnb = 1.0 if base <= 0 else float(np.log(base))
        if T == "EXP":
            return _Lambda(lambda x: torch.exp(lnb * (sh + sc * x)), "exp"), in_channels
        return _Lambda(lambda x: torch.log(sh + sc * x) / lnb, "log"), in_channels
    if T == "THRESHOLD":
        th = float(_one(_sub(layer, "threshold_param"), "threshold", 0.0))
        return _Lambda(lambda x: (x > th).to(x.dtype), f"threshold({th})"), in_channels
    if T == "TILE":
        p = _sub(layer, "tile_param")
        axis, tiles = int(_one(p, "axis", 1)), int(_one(p, "tiles"))

        def tile(x):
            reps = [1] * x.dim()
            reps[axis] = tiles
            return x.repeat(*reps)

        return _Lambda(tile, f"tile(axis={axis}, tiles={tiles})"), None
    if T == "SPLIT":
        return None, in_channels
    raise CaffeConversionException(f"unsupported caffe layer type {ltype!r} (pass a converter through customized_converters)")


_SKIP = {"DATA", "IMAGEDATA", "HDF5DATA", "HDF5OUTPUT", "WINDOWDATA", "MEMORYDATA", "DUMMYDATA", "ANNOTATEDDATA", "ACCURACY", "SILENCE",
         "EUCLIDEANLOSS", "HINGELOSS", "INFOGAINLOSS", "MULTINOMIALLOGISTICLOSS", "CONTRASTIVELOSS"}


class CaffeGraph(nn.Module):
    """The converted network: nodes run in definition order over a dict of named blobs."""

    def __init__(self, name, inputs, nodes, outputs):
        super().__init__()
        self.net_name, self.input_names, self.output_names = name, list(inputs), list(outputs)
        self.plan = []
        self.layers = nn.ModuleDict()
        for lname, module, bottoms, tops in nodes:
            key = lname.replace(".", "_").replace("/", "_")
            base, k = key, 1
            while key in self.layers:
                key, k = f"{base}_{k}", k + 1
            self.layers[key] = module
            self.plan.append((key, lname, list(bottoms), list(tops)))

    def forward(self, *xs):
        if len(xs) == 1 and isinstance(xs[0], (tuple, list)):
            xs = tuple(xs[0])
        if len(xs) != len(self.input_names):
            raise ValueError(f"{self.net_name}: expected {len(self.input_names)} inputs {self.input_names}, got {len(xs)}")
        blobs = dict(zip(self.input_names, xs))
        for key, _, bottoms, tops in self.plan:
            m = self.layers[key]
            y = m(*[blobs[b] for b in bottoms])
            if isinstance(y, (tuple, list)):
                for t, v in zip(tops, y):
                    blobs[t] = v
            else:
                blobs[tops[0]] = y
        outs = [blobs[o] for o in self.output_names]
        return outs[0] if len(outs) == 1 else tuple(outs)


class _Slice(nn.Module):
    def __init__(self, axis, points, n):
        super().__init__()
        self.axis, self.points, self.n = axis, points, n

    def forward(self, x):
        if self.points:
            sizes, prev = [], 0
            for p in self.points:
                sizes.append(p - prev)
                prev = p
            sizes.append(x.shape[self.axis] - prev)
            return torch.split(x, sizes, self.axis)
        return torch.chunk(x, self.n, self.axis)


class CaffeLoader:
    def __init__(self, def_path, model_path=None, match_all=True, customized_converters=None):
        self.def_path, self.model_path, self.match_all = def_path, model_path, match_all
        self.custom = {k.upper(): v for k, v in (customized_converters or {}).items()}

    # ---- reading ----------------------------------------------------------------------------------------------------------
    @staticmethod
    def read_definition(path):
        with open(path, "r") as f:
            return pw.parse_text(f.read())

    @staticmethod
    def read_weights(path):
        """{layer name: [tensors]} from a binary NetParameter (V2 `layer` and V1 `layers` entries)."""
        with open(path, "rb") as f:
            net = pw.decode(f.read(), cs.NET)
        out = {}
        for lay in net.get("layer", []) + net.get("layers", []):
            bl = [_blob_array(b) for b in lay.get("blobs", [])]
            if bl:
                out[_one(lay, "name")] = bl
        return out

    @staticmethod
    def _layers_of(net):
        layers = []
        for lay in net.get("layer", []):
            layers.append((str(_one(lay, "type", "")), lay))
        for lay in net.get("layers", []):
            t = _one(lay, "type", 0)
            layers.append((cs.V1_TYPES.get(t, str(t)) if isinstance(t, int) else str(t), lay))
        return layers

    @staticmethod
    def _train_only(lay):
        ph = lambda r: {"TRAIN": 0, "TEST": 1}.get(_one(r, "phase"), _one(r, "phase"))
        inc = [ph(r) for r in lay.get("include", [])]
        return bool(inc) and all(p == 0 for p in inc)

    # ---- conversion -------------------------------------------------------------------------------------------------------
    def create_module(self):
        net = self.read_definition(self.def_path)
        weights = self.read_weights(self.model_path) if self.model_path else {}
        inputs, channels = [], {}
        dims = [int(d) for d in net.get("input_dim", [])]
        shapes = [[int(d) for d in s.get("dim", [])] for s in net.get("input_shape", [])]
        for i, name in enumerate(net.get("input", [])):
            inputs.append(name)
            shp = shapes[i] if i < len(shapes) else dims[4 * i:4 * i + 4]
            channels[name] = shp[1] if len(shp) > 1 else None
        nodes, produced, consumed, used_weights = [], set(inputs), set(), set()
        for ltype, lay in self._layers_of(net):
            T = ltype.upper().replace("_", "")
            name = str(_one(lay, "name", ltype))
            bottoms, tops = [str(b) for b in lay.get("bottom", [])], [str(t) for t in lay.get("top", [])]
            if T == "INPUT":
                shp = [int(d) for d in _sub(_sub(lay, "input_param"), "shape").get("dim", [])]
                for t in tops:
                    inputs.append(t)
                    produced.add(t)
                    channels[t] = shp[1] if len(shp) > 1 else None
                continue
            if T in _SKIP or self._train_only(lay):
                if not bottoms:                       # a data layer: its first top is a network input
                    if tops and tops[0] not in produced:
                        inputs.append(tops[0])
                        produced.add(tops[0])
                        channels[tops[0]] = None
                continue
            if T.endswith("LOSS") and len(bottoms) > 1:
                bottoms = bottoms[:1]                 # (prediction, label) -> prediction
            blobs = weights.get(name, [])
            if blobs:
                used_weights.add(name)
            cin = channels.get(bottoms[0]) if bottoms else None
            if T in self.custom:
                module, cout = self.custom[T](lay, blobs), None
            elif T == "SLICE":
                p = _sub(lay, "slice_param")
                module = _Slice(int(_one(p, "axis", _one(p, "slice_dim", 1))), [int(x) for x in p.get("slice_point", [])], len(tops))
                cout = None
            else:
                module, cout = _convert(ltype, lay, blobs, cin)
            if module is None:                        # Split: every top aliases the bottom
                module = _Lambda(lambda x, n=len(tops): x if n == 1 else tuple([x] * n), "split")
            if self.match_all and self.model_path and not blobs and any(True for _ in module.parameters()):
                raise CaffeConversionException(f"layer {name!r} has parameters but no weights in {self.model_path} (match_all=True)")
            nodes.append((name, module, bottoms, tops))
            consumed.update(bottoms)
            for t in tops:
                produced.add(t)
                channels[t] = cout
        outputs = []
        for _, _, bottoms, tops in nodes:
            for t in tops:
                if t not in outputs:
                    outputs.append(t)
        last_writer = {}
        for i, (_, _, _, tops) in enumerate(nodes):
            for t in tops:
                last_writer[t] = i
        finals = []
        for t in outputs:
            later_readers = [i for i, (_, _, b, _) in enumerate(nodes) if t in b and i > last_writer[t]]
            if not later_readers:
                finals.append(t)
        graph = CaffeGraph(str(_one(net, "name", "caffe_net")), inputs, nodes, finals)
        graph.unused_weights = sorted(set(weights) - used_weights)
        return graph

    # ---- copy weights into an existing model ---------------------------------------------------------------------------------
    @staticmethod
    def load(model, def_path, model_path, match_all=True, customized_converters=None):
        """Copy the caffemodel's weights into `model` (an nn.Module) by sub-module name (reference: CaffeLoader.load, :684-700).
        Returns the model."""
        weights = CaffeLoader.read_weights(model_path)
        mods = dict(model.named_modules())
        for name, blobs in weights.items():
            m = mods.get(name) or mods.get(name.replace("/", "_").replace(".", "_"))
            if m is None:
                continue
            params = [p for p in (getattr(m, "weight", None), getattr(m, "bias", None)) if p is not None]
            with torch.no_grad():
                for p, b in zip(params, blobs):
                    if p.numel() != b.numel():
                        raise CaffeConversionException(f"layer {name}: weight size {b.numel()} does not match the model's {p.numel()}")
                    p.copy_(b.reshape(p.shape))
        if match_all:
            missing = [n for n, m in mods.items() if n and any(True for _ in m.parameters(recurse=False)) and n not in weights
                       and n.replace("_", "/") not in weights]
            if missing:
                raise CaffeConversionException(f"modules without weights in {model_path}: {missing} (match_all=True)")
        return model

    @staticmethod
    def load_caffe(def_path, model_path=None, match_all=True, customized_converters=None):
        """prototxt (+ caffemodel) -> `CaffeGraph` (reference: CaffeLoader.loadCaffe / Net.loadCaffe)."""
        return CaffeLoader(def_path, model_path, match_all, customized_converters).create_module()


def load_caffe(def_path, model_path=None, match_all=True, customized_converters=None):
    return CaffeLoader.load_caffe(def_path, model_path, match_all, customized_converters)


def save_caffemodel(path, layers, name="net"):
    """Write a binary NetParameter holding `layers` = [(layer name, type string, [weight tensors])] -- the exporter counterpart
    of `read_weights` (used by tests and by `Net.save` round trips of imported models)."""
    msg = {"name": name, "layer": []}
    for lname, ltype, blobs in layers:
        msg["layer"].append({"name": lname, "type": ltype,
                             "blobs": [{"shape": {"dim": list(b.shape)}, "data": b.detach().float().reshape(-1).numpy()} for b in blobs]})
    with open(path, "wb") as f:
        f.write(pw.encode(msg, cs.NET))
