"""Load models saved by BigDL / Analytics Zoo (`Module.saveModule`, `ZooModel.saveModel`, `KerasNet.saveModule`) as torch modules.

Reference: `Net.load_bigdl` / `Net.loadBigDL` (pyzoo/zoo/pipeline/api/net/net_load.py:60-75, zoo/.../pipeline/api/Net.scala:140-170)
hand the file to BigDL's `ModuleLoader`, which rebuilds JVM layers by reflection.  Here the protobuf is decoded without generated
classes (:mod:`zoo.util.protowire` + `bigdl_schema`), every `moduleType` is mapped to a torch module by a converter, and the
parameters are copied from the file's global storage.  Containers: `Sequential`, `StaticGraph` / `Graph` (executed in topological
order over `preModules`), and the Keras-style zoo layers (`com.intel.analytics.zoo.pipeline.api.keras.*`), which wrap their
BigDL implementation as their only sub-module.  Unknown layer types raise `BigDLConversionException` naming the layer; register
more with `register_bigdl_layer`.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ...util import protowire as pw
from . import bigdl_schema as S


class BigDLConversionException(Exception):
    pass


# ---------------------------------------------------------------------------------------------------------------- parsing --
def _attrs(mod):
    return {pw.one(e, "key"): pw.one(e, "value") for e in mod.get("attr", [])}


def _mod_name(mod):
    """BigDL's getName(): the given name, or class name + the per-object postfix."""
    n = pw.one(mod, "name", "")
    return n if n else pw.one(mod, "moduleType", "").rsplit(".", 1)[-1] + pw.one(mod, "namePostfix", "")


def _attr_value(av, storages=None):
    """AttrValue -> python value (ints, floats, strings, bools, lists, arrays, tensors)."""
    if av is None:
        return None
    for key in ("stringValue", "boolValue", "floatValue", "doubleValue", "int64Value", "int32Value", "dataFormatValue"):
        if key in av:
            return av[key][-1]
    if "tensorValue" in av:
        return _tensor(av["tensorValue"][-1], storages or {})
    if "arrayValue" in av:
        arr = av["arrayValue"][-1]
        for key in ("i32", "i64", "str", "boolean"):
            if key in arr:
                return list(arr[key])
        for key in ("flt", "dbl"):
            if key in arr:
                return list(np.concatenate([np.atleast_1d(np.asarray(p)) for p in arr[key]]))
        if "tensor" in arr:
            return [_tensor(t, storages or {}) for t in arr["tensor"]]
        return []
    if "shape" in av:
        return list(av["shape"][-1].get("shapeValue", []))
    if "nameAttrListValue" in av:
        return av["nameAttrListValue"][-1]
    if "bigDLModuleValue" in av:
        return av["bigDLModuleValue"][-1]
    dt = pw.one(av, "dataType", 0)
    return 0 if dt in (S.DT_INT32, S.DT_INT64) else (False if dt == S.DT_BOOL else None)      # proto3 omits zero values


def _storage_data(st):
    if st is None:
        return None
    if st.get("float_data"):
        return pw.floats(st, "float_data")
    if st.get("double_data"):
        return np.concatenate([np.atleast_1d(np.asarray(p, dtype=np.float64)) for p in st["double_data"]])
    for key in ("int_data", "long_data"):
        if st.get(key):
            return np.asarray(st[key], dtype=np.int64)
    return None


def _collect_storages(root):
    """storage id -> flat ndarray, from the root module's `global_storage` attribute (and any tensor that carries data inline)."""
    out = {}
    gs = _attrs(root).get("global_storage")
    if gs is not None and "nameAttrListValue" in gs:
        for e in gs["nameAttrListValue"][-1].get("attr", []):
            val = pw.one(e, "value") or {}
            for t in val.get("tensorValue", []):
                st = pw.one(t, "storage")
                data = _storage_data(st)
                if data is not None:
                    out[pw.one(st, "id", 0)] = data
    return out


def _tensor(t, storages):
    """BigDLTensor -> torch tensor (None for an empty tensor)."""
    if t is None:
        return None
    size = [int(x) for x in t.get("size", [])]
    st = pw.one(t, "storage")
    if st is None or (not size and not pw.one(t, "isScalar", False)):
        return None
    data = _storage_data(st)
    if data is None:
        data = storages.get(pw.one(st, "id", 0))
    if data is None:
        raise BigDLConversionException(f"tensor {pw.one(t, 'id')} refers to storage {pw.one(st, 'id')}, which is not in the file")
    flat = torch.from_numpy(np.ascontiguousarray(data))
    stride = [int(x) for x in t.get("stride", [])]
    off = max(int(pw.one(t, "offset", 1)) - 1, 0)                      # storage offsets are 1-based
    if not size:
        return flat[off].clone()
    if not stride:
        stride, acc = [], 1
        for s in reversed(size):
            stride.insert(0, acc)
            acc *= s
    return torch.as_strided(flat, size, stride, off).clone().float()


# ------------------------------------------------------------------------------------------------------------- torch side --
class _Lambda(nn.Module):
    def __init__(self, fn, desc=""):
        super().__init__()
        self.fn, self.desc = fn, desc

    def forward(self, *xs):
        return self.fn(*xs)

    def extra_repr(self):
        return self.desc


class _Reshape(nn.Module):
    """BigDL Reshape(size, batchMode): batchMode true keeps dim 0, false never does, unset decides by the element count."""

    def __init__(self, size, batch_mode=None):
        super().__init__()
        self.size, self.batch_mode = [int(s) for s in size], batch_mode

    def forward(self, x):
        n = int(np.prod(self.size))
        batch = self.batch_mode if self.batch_mode is not None else not (x.numel() == n and x.shape[0] != 1)
        return x.reshape(x.shape[0], *self.size) if batch else x.reshape(*self.size)


class BigDLGraph(nn.Module):
    """A `StaticGraph`: nodes run in topological order; a node with several predecessors receives them as separate arguments
    (table inputs of BigDL), graph inputs are fed in `input_names` order."""

    def __init__(self, nodes, edges, input_names, output_names):
        super().__init__()
        self.nodes = nn.ModuleDict({self._key(k): m for k, m in nodes.items()})
        self.edges = {k: list(v) for k, v in edges.items()}
        self.input_names, self.output_names = list(input_names), list(output_names)
        order, seen = [], set()

        def visit(n):
            if n in seen:
                return
            seen.add(n)
            for p in self.edges.get(n, []):
                visit(p)
            order.append(n)

        for o in self.output_names:
            visit(o)
        self.order = order

    @staticmethod
    def _key(name):
        return name.replace(".", "_dot_")

    def forward(self, *xs):
        vals = {}
        feeds = dict(zip(self.input_names, xs))
        for n in self.order:
            pre = self.edges.get(n, [])
            args = [vals[p] for p in pre] if pre else [feeds[n]] if n in feeds else list(xs[:1])
            vals[n] = self.nodes[self._key(n)](*args)
        outs = [vals[o] for o in self.output_names]
        return outs[0] if len(outs) == 1 else tuple(outs)


_CONVERTERS = {}


def register_bigdl_layer(*type_names):
    """Decorator: `fn(ctx) -> nn.Module` for the given BigDL class names (the last path component, e.g. "Linear")."""
    def deco(fn):
        for t in type_names:
            _CONVERTERS[t] = fn
        return fn
    return deco


class _Ctx:
    """What a converter sees: attributes as python values, parameters as tensors, converted sub-modules."""

    def __init__(self, mod, storages, loader):
        self.mod, self.storages, self.loader = mod, storages, loader
        self.name = _mod_name(mod)
        self.type = pw.one(mod, "moduleType", "")
        self._a = _attrs(mod)

    def attr(self, key, default=None):
        if key not in self._a:
            return default
        v = _attr_value(self._a[key], self.storages)
        return default if v is None else v

    # BigDL 0.5 / 0.6 files carry `weight` and `bias`; later versions write the layer's parameters() list instead
    # (`hasParameters` + `parameters`: weight first, bias second)
    @property
    def weight(self):
        t = pw.one(self.mod, "weight")
        if t is None and self.mod.get("parameters"):
            t = self.mod["parameters"][0]
        return _tensor(t, self.storages)

    @property
    def bias(self):
        t = pw.one(self.mod, "bias")
        if t is None and pw.one(self.mod, "weight") is None and len(self.mod.get("parameters", [])) > 1:
            t = self.mod["parameters"][1]
        return _tensor(t, self.storages)

    @property
    def parameters(self):
        return [_tensor(t, self.storages) for t in self.mod.get("parameters", [])]

    def children(self):
        return [self.loader.convert(m) for m in self.mod.get("subModules", [])]


def _pair(ctx, a, b, default=0):
    return int(ctx.attr(a, default)), int(ctx.attr(b, default))


# -------------------------------------------------------------------------------------------------------------- converters --
@register_bigdl_layer("Sequential")
def _seq(ctx):
    return nn.Sequential(*ctx.children())


@register_bigdl_layer("StaticGraph", "Graph", "DynamicGraph", "Model")
def _graph(ctx):
    subs = ctx.mod.get("subModules", [])
    if ctx.type.startswith("com.intel.analytics.zoo.pipeline.api.keras") and len(subs) == 1:
        return ctx.loader.convert(subs[0])                       # keras-style Model: its labor is the graph
    nodes, edges = {}, {}
    for m in subs:
        n = _mod_name(m)
        nodes[n] = ctx.loader.convert(m)
        edges[n] = list(m.get("preModules", []))
    names = set(nodes)
    ins = ctx.attr("inputNames") or [n for n in nodes if not edges[n]]
    outs = ctx.attr("outputNames")
    if not outs:
        used = {p for ps in edges.values() for p in ps}
        outs = [n for n in nodes if n not in used]
    for n, ps in edges.items():
        for q in ps:
            if q not in names:
                raise BigDLConversionException(f"graph node {n} depends on {q}, which is not in the graph")
    return BigDLGraph(nodes, edges, ins, outs)


@register_bigdl_layer("Input", "Identity", "InputLayer")
def _identity(ctx):
    return nn.Identity()


@register_bigdl_layer("Linear")
def _linear(ctx):
    w, b = ctx.weight, ctx.bias
    lin = nn.Linear(int(ctx.attr("inputSize", w.shape[1])), int(ctx.attr("outputSize", w.shape[0])), bias=b is not None)
    with torch.no_grad():
        lin.weight.copy_(w.reshape(lin.weight.shape))
        if b is not None:
            lin.bias.copy_(b.reshape(-1))
    return lin


@register_bigdl_layer("SpatialConvolution", "SpatialShareConvolution", "SpatialDilatedConvolution")
def _conv(ctx):
    if int(ctx.attr("format", 0)) != 0:
        raise BigDLConversionException(f"{ctx.name}: NHWC convolutions are not supported")
    cin, cout, g = int(ctx.attr("nInputPlane")), int(ctx.attr("nOutputPlane")), int(ctx.attr("nGroup", 1))
    kw, kh = _pair(ctx, "kernelW", "kernelH", 1) if ctx.attr("kernelW") is not None else _pair(ctx, "kW", "kH", 1)
    sw, sh = _pair(ctx, "strideW", "strideH", 1) if ctx.attr("strideW") is not None else _pair(ctx, "dW", "dH", 1)
    pw_, ph = _pair(ctx, "padW", "padH", 0)
    dil = (int(ctx.attr("dilationH", 1)), int(ctx.attr("dilationW", 1)))
    if pw_ < 0 or ph < 0:                                        # -1: TensorFlow-style SAME padding
        if sw != 1 or sh != 1:
            raise BigDLConversionException(f"{ctx.name}: SAME padding with stride > 1 is not supported")
        ph, pw_ = dil[0] * (kh - 1) // 2, dil[1] * (kw - 1) // 2
    w, b = ctx.weight, ctx.bias
    conv = nn.Conv2d(cin, cout, (kh, kw), (sh, sw), (ph, pw_), dil, g, bias=b is not None)
    with torch.no_grad():
        conv.weight.copy_(w.reshape(conv.weight.shape))        # stored as [group, out / group, in / group, kH, kW]
        if b is not None:
            conv.bias.copy_(b.reshape(-1))
    return conv


def _pool(ctx, cls):
    kw, kh = _pair(ctx, "kW", "kH", 1)
    dw, dh = _pair(ctx, "dW", "dH", 1)
    pw_, ph = _pair(ctx, "padW", "padH", 0)
    if int(ctx.attr("format", 0)) != 0:
        raise BigDLConversionException(f"{ctx.name}: NHWC pooling is not supported")
    kwargs = dict(ceil_mode=bool(ctx.attr("ceil_mode", ctx.attr("ceilMode", False))))
    if cls is nn.AvgPool2d:
        if ctx.attr("globalPooling", False):
            return nn.AdaptiveAvgPool2d(1)
        kwargs["count_include_pad"] = bool(ctx.attr("countIncludePad", True))
    return cls((kh, kw), (dh, dw), (max(ph, 0), max(pw_, 0)), **kwargs)


@register_bigdl_layer("SpatialMaxPooling")
def _maxpool(ctx):
    return _pool(ctx, nn.MaxPool2d)


@register_bigdl_layer("SpatialAveragePooling")
def _avgpool(ctx):
    return _pool(ctx, nn.AvgPool2d)


def _bn(ctx, cls):
    n = int(ctx.attr("nOutput"))
    affine = bool(ctx.attr("affine", True))
    bn = cls(n, eps=float(ctx.attr("eps", 1e-5)), momentum=float(ctx.attr("momentum", 0.1)), affine=affine)
    with torch.no_grad():
        if affine and ctx.weight is not None:
            bn.weight.copy_(ctx.weight.reshape(-1))
            bn.bias.copy_(ctx.bias.reshape(-1))
        mean, var = ctx.attr("runningMean"), ctx.attr("runningVar")
        if torch.is_tensor(mean):
            bn.running_mean.copy_(mean.reshape(-1))
        if torch.is_tensor(var):
            bn.running_var.copy_(var.reshape(-1))
    return bn


@register_bigdl_layer("BatchNormalization")
def _bn1(ctx):
    return _bn(ctx, nn.BatchNorm1d)


@register_bigdl_layer("SpatialBatchNormalization")
def _bn2(ctx):
    return _bn(ctx, nn.BatchNorm2d)


for _name, _cls in (("ReLU", nn.ReLU), ("ReLU6", nn.ReLU6), ("Tanh", nn.Tanh), ("Sigmoid", nn.Sigmoid), ("SoftPlus", nn.Softplus), ("Abs", None),
                    ("SoftSign", nn.Softsign), ("HardTanh", nn.Hardtanh), ("Exp", None), ("Log", None), ("Sqrt", None), ("Square", None)):
    def _make(cls, name):
        fns = {"Abs": torch.abs, "Exp": torch.exp, "Log": torch.log, "Sqrt": torch.sqrt, "Square": torch.square}
        return (lambda ctx: cls()) if cls is not None else (lambda ctx: _Lambda(fns[name], name))
    register_bigdl_layer(_name)(_make(_cls, _name))


@register_bigdl_layer("SoftMax")
def _softmax(ctx):
    return nn.Softmax(dim=-1)


@register_bigdl_layer("LogSoftMax")
def _logsoftmax(ctx):
    return nn.LogSoftmax(dim=-1)


@register_bigdl_layer("ELU")
def _elu(ctx):
    return nn.ELU(float(ctx.attr("alpha", 1.0)))


@register_bigdl_layer("LeakyReLU")
def _leaky(ctx):
    return nn.LeakyReLU(float(ctx.attr("negval", 0.01)))


@register_bigdl_layer("Dropout", "SpatialDropout2D", "GaussianDropout")
def _dropout(ctx):
    return nn.Dropout(float(ctx.attr("initP", ctx.attr("p", 0.5))))


@register_bigdl_layer("Reshape")
def _reshape(ctx):
    mode = int(ctx.attr("batchMode", 0))
    return _Reshape(ctx.attr("size"), {0: None, 1: True, 2: False}.get(mode))


@register_bigdl_layer("InferReshape")
def _infer_reshape(ctx):
    size = [int(v) for v in ctx.attr("size")]
    batch = bool(ctx.attr("batchMode", False))

    def fn(x):
        tgt = [x.shape[i + (1 if batch else 0)] if v == 0 else v for i, v in enumerate(size)]      # 0 copies the input dimension
        return x.reshape(x.shape[0], *tgt) if batch else x.reshape(*tgt)
    return _Lambda(fn, f"InferReshape({size}, batch={batch})")


@register_bigdl_layer("View")
def _view(ctx):
    sizes = [int(v) for v in ctx.attr("sizes")]
    return _Lambda(lambda x: x.reshape(-1, *sizes) if x.numel() != int(np.prod(sizes)) else x.reshape(*sizes), f"View({sizes})")


@register_bigdl_layer("Flatten")
def _flatten(ctx):
    return nn.Flatten()


@register_bigdl_layer("Squeeze")
def _squeeze(ctx):
    dims = ctx.attr("dims") or ([ctx.attr("dim")] if ctx.attr("dim") is not None else None)
    batch = bool(ctx.attr("batchMode", False))
    if not dims:
        return _Lambda(lambda x: x.squeeze(), "Squeeze()")
    ds = sorted((int(d) - 1 + (1 if batch else 0) for d in dims), reverse=True)             # BigDL dimensions are 1-based

    def fn(x):
        for d in ds:
            x = x.squeeze(d)
        return x
    return _Lambda(fn, f"Squeeze({ds})")


@register_bigdl_layer("Unsqueeze")
def _unsqueeze(ctx):
    pos = int(ctx.attr("pos")) - 1
    return _Lambda(lambda x: x.unsqueeze(pos), f"Unsqueeze({pos})")


@register_bigdl_layer("Transpose")
def _transpose(ctx):
    perm = [int(v) - 1 for v in ctx.attr("permutations")]
    pairs = list(zip(perm[0::2], perm[1::2]))

    def fn(x):
        for a, b in pairs:
            x = x.transpose(a, b)
        return x
    return _Lambda(fn, f"Transpose({pairs})")


@register_bigdl_layer("CAddTable")
def _cadd(ctx):
    return _Lambda(lambda *xs: sum(xs[1:], xs[0]) if len(xs) > 1 else sum(xs[0][1:], xs[0][0]), "CAddTable")


@register_bigdl_layer("CMulTable")
def _cmul(ctx):
    def fn(*xs):
        xs = xs if len(xs) > 1 else tuple(xs[0])
        out = xs[0]
        for t in xs[1:]:
            out = out * t
        return out
    return _Lambda(fn, "CMulTable")


@register_bigdl_layer("JoinTable")
def _join(ctx):
    dim, nin = int(ctx.attr("dimension")), int(ctx.attr("nInputDims", 0))

    def fn(*xs):
        xs = xs if len(xs) > 1 else tuple(xs[0])
        d = dim - 1 + (1 if nin > 0 and xs[0].dim() > nin else 0)
        return torch.cat(xs, d)
    return _Lambda(fn, f"JoinTable({dim})")


@register_bigdl_layer("LookupTable")
def _lookup(ctx):
    w = ctx.weight
    emb = nn.Embedding(w.shape[0], w.shape[1])
    with torch.no_grad():
        emb.weight.copy_(w)
    return nn.Sequential(_Lambda(lambda x: x.long() - 1, "one-based ids"), emb)


@register_bigdl_layer("MulConstant")
def _mulc(ctx):
    c = float(ctx.attr("scalar", ctx.attr("constant", 1.0)))
    return _Lambda(lambda x: x * c, f"MulConstant({c})")


@register_bigdl_layer("AddConstant")
def _addc(ctx):
    c = float(ctx.attr("constant_scalar", ctx.attr("constant", 0.0)))
    return _Lambda(lambda x: x + c, f"AddConstant({c})")


@register_bigdl_layer("TimeDistributed")
def _time_distributed(ctx):
    inner = ctx.children()[0]

    def fn(x):
        y = inner(x.reshape(-1, *x.shape[2:]))
        return y.reshape(x.shape[0], x.shape[1], *y.shape[1:])
    m = _Lambda(fn, "TimeDistributed")
    m.inner = inner
    return m


# ------------------------------------------------------------------------------------------------------------------ loader --
WEIGHT_MAGIC = 3721


def read_weight_file(path_or_bytes):
    """The separate weight file of `saveModule(path, weightPath)` -> {storage id: flat float array}.

    Layout (big-endian, written through a Java DataOutputStream): int magic 3721, int n, then n records of (int storage id, long
    element count, that many floats), then an int length and an MD5 digest of the payload."""
    import struct

    buf = path_or_bytes
    if not isinstance(buf, (bytes, bytearray, memoryview)):
        from ...util import filesystem as zfs

        with zfs.open_file(path_or_bytes, "rb") as fh:
            buf = fh.read()
    magic, n = struct.unpack_from(">ii", buf, 0)
    if magic != WEIGHT_MAGIC:
        raise BigDLConversionException(f"not a BigDL weight file (magic {magic}, expected {WEIGHT_MAGIC})")
    out, pos = {}, 8
    for _ in range(n):
        sid, count = struct.unpack_from(">iq", buf, pos)
        pos += 12
        out[sid] = np.frombuffer(buf, dtype=">f4", count=count, offset=pos).astype(np.float32)
        pos += 4 * count
    return out


class BigDLLoader:
    def __init__(self, path_or_bytes, weight_path=None):
        if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
            buf = bytes(path_or_bytes)
        else:
            from ...util import filesystem as zfs

            with zfs.open_file(path_or_bytes, "rb") as fh:
                buf = fh.read()
        self.root = pw.decode(buf, S.MODULE)
        if "moduleType" not in self.root:
            raise BigDLConversionException("not a BigDL model file (no moduleType in the root message)")
        self.storages = _collect_storages(self.root)
        if weight_path is not None:                   # definition and weights saved apart: storages are looked up by id in the weight file
            self.storages.update(read_weight_file(weight_path))

    def convert(self, mod):
        mtype = pw.one(mod, "moduleType", "")
        short = mtype.rsplit(".", 1)[-1]
        subs = mod.get("subModules", [])
        conv = _CONVERTERS.get(short)
        keras = mtype.startswith("com.intel.analytics.zoo.pipeline.api.keras") or pw.one(_attrs(mod).get("is_keras_module") or {}, "boolValue", False)
        if keras and len(subs) == 1 and short not in ("StaticGraph", "Graph"):
            # a Keras-style layer serialises its BigDL implementation ("labor") as its only sub-module
            out = self.convert(subs[0])
        elif conv is not None:
            out = conv(_Ctx(mod, self.storages, self))
        else:
            raise BigDLConversionException(f"no converter for BigDL layer type {mtype} (layer {_mod_name(mod)!r}); "
                                           "register one with zoo.models.bigdl.register_bigdl_layer")
        out._bigdl_name = _mod_name(mod)
        return out

    def load(self):
        net = self.convert(self.root)
        return net.train(bool(pw.one(self.root, "train", False)))


def load_bigdl(path_or_bytes, weight_path=None):
    """BigDL / zoo model file (+ the separate weight file of `saveModule(path, weightPath)`) -> torch module (eval mode unless the
    file says it was saved in training mode)."""
    return BigDLLoader(path_or_bytes, weight_path).load()
