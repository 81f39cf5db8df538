"""Write a torch model as a BigDL model file (the format of `Module.saveModule` / `ZooModel.saveModel`; reference:
`KerasNet.saveModule`, zoo/.../pipeline/api/keras/models/Topology.scala, and `Net.saveToKeras2 / saveToTf` as sibling exporters,
zoo/.../pipeline/api/Net.scala:300-446).  Feed-forward models made of the layer types below are supported; the result loads in
BigDL / Analytics Zoo (`Module.loadModule`, `Net.load_bigdl`) and in this framework's `load_bigdl`."""
from __future__ import annotations

import itertools

import numpy as np
import torch.nn as nn

from ...util import protowire as pw
from . import bigdl_schema as S

_NN = "com.intel.analytics.bigdl.nn."


class _Writer:
    def __init__(self):
        self.ids = itertools.count(1000)
        self.storage_entries = []
        self.anon = itertools.count(1)

    def tensor(self, t):
        """Reference to a parameter: shape + storage id; the data goes to the global storage once."""
        arr = np.ascontiguousarray(t.detach().float().cpu().numpy())
        tid, sid = next(self.ids), next(self.ids)
        size = list(arr.shape)
        stride, acc = [], 1
        for s in reversed(size):
            stride.insert(0, acc)
            acc *= s
        head = {"datatype": S.DT_FLOAT, "size": size, "stride": stride, "offset": 1, "dimension": len(size), "nElements": int(arr.size), "id": tid}
        self.storage_entries.append({"key": str(tid), "value": {"dataType": S.DT_TENSOR, "tensorValue": dict(
            head, storage={"datatype": S.DT_FLOAT, "float_data": arr.reshape(-1), "id": sid})}})
        return dict(head, storage={"datatype": S.DT_FLOAT, "id": sid})

    @staticmethod
    def a_int(v):
        return {"dataType": S.DT_INT32, "int32Value": int(v)}

    @staticmethod
    def a_bool(v):
        return {"dataType": S.DT_BOOL, "boolValue": bool(v)}

    @staticmethod
    def a_double(v):
        return {"dataType": S.DT_DOUBLE, "doubleValue": float(v)}

    @staticmethod
    def a_ints(vs):
        return {"dataType": S.DT_ARRAY, "arrayValue": {"size": len(vs), "datatype": S.DT_INT32, "i32": [int(v) for v in vs]}}

    def module(self, short, name=None, attrs=None, subs=None, weight=None, bias=None):
        postfix = "%08x" % next(self.anon)
        m = {"subModules": subs or []}
        if name:
            m["name"] = name
        if weight is not None:
            m["weight"] = weight
        if bias is not None:
            m["bias"] = bias
        m["moduleType"] = _NN + short
        m["attr"] = [{"key": k, "value": v} for k, v in (attrs or {}).items()]
        m["version"] = "0.5.0"
        m["train"] = False
        m["namePostfix"] = postfix
        m["id"] = next(self.ids)
        # field order of `encode` follows this dict; protobuf readers accept any order
        return {k: m[k] for k in ("name", "subModules", "weight", "bias", "moduleType", "attr", "version", "train", "namePostfix", "id") if k in m}

    def convert(self, mod, name=None):
        two = lambda v: (v, v) if isinstance(v, int) else tuple(v)
        if isinstance(mod, nn.Sequential):
            return self.module("Sequential", name, subs=[self.convert(c, n) for n, c in mod.named_children()])
        if isinstance(mod, nn.Linear):
            return self.module("Linear", name, {"inputSize": self.a_int(mod.in_features), "outputSize": self.a_int(mod.out_features),
                                                "withBias": self.a_bool(mod.bias is not None)},
                               weight=self.tensor(mod.weight), bias=None if mod.bias is None else self.tensor(mod.bias))
        if isinstance(mod, nn.Conv2d):
            if mod.padding_mode != "zeros" or isinstance(mod.padding, str):
                raise NotImplementedError("save_bigdl: Conv2d needs numeric zero padding")
            kh, kw = mod.kernel_size
            g = mod.groups
            w = mod.weight.reshape(g, mod.out_channels // g, mod.in_channels // g, kh, kw)
            short = "SpatialConvolution" if mod.dilation == (1, 1) else "SpatialDilatedConvolution"
            at = {"nInputPlane": self.a_int(mod.in_channels), "nOutputPlane": self.a_int(mod.out_channels), "kernelW": self.a_int(kw),
                  "kernelH": self.a_int(kh), "strideW": self.a_int(mod.stride[1]), "strideH": self.a_int(mod.stride[0]),
                  "padW": self.a_int(mod.padding[1]), "padH": self.a_int(mod.padding[0]), "nGroup": self.a_int(g),
                  "withBias": self.a_bool(mod.bias is not None), "format": {"dataType": S.DT_DATA_FORMAT, "dataFormatValue": 0}}
            if short == "SpatialDilatedConvolution":
                at.update({"dilationW": self.a_int(mod.dilation[1]), "dilationH": self.a_int(mod.dilation[0])})
            return self.module(short, name, at, weight=self.tensor(w), bias=None if mod.bias is None else self.tensor(mod.bias))
        if isinstance(mod, (nn.MaxPool2d, nn.AvgPool2d)):
            (kh, kw), (ph, pw_) = two(mod.kernel_size), two(mod.padding)
            dh, dw = two(mod.stride if mod.stride is not None else mod.kernel_size)
            at = {"kW": self.a_int(kw), "kH": self.a_int(kh), "dW": self.a_int(dw), "dH": self.a_int(dh), "padW": self.a_int(pw_), "padH": self.a_int(ph),
                  "format": {"dataType": S.DT_DATA_FORMAT, "dataFormatValue": 0}, "ceil_mode": self.a_bool(mod.ceil_mode)}
            if isinstance(mod, nn.AvgPool2d):
                at.update({"ceilMode": self.a_bool(mod.ceil_mode), "countIncludePad": self.a_bool(mod.count_include_pad), "globalPooling": self.a_bool(False)})
            return self.module("SpatialMaxPooling" if isinstance(mod, nn.MaxPool2d) else "SpatialAveragePooling", name, at)
        if isinstance(mod, (nn.BatchNorm1d, nn.BatchNorm2d)):
            at = {"nOutput": self.a_int(mod.num_features), "eps": self.a_double(mod.eps), "momentum": self.a_double(mod.momentum or 0.1),
                  "affine": self.a_bool(mod.affine),
                  "runningMean": {"dataType": S.DT_TENSOR, "tensorValue": self.tensor(mod.running_mean)},
                  "runningVar": {"dataType": S.DT_TENSOR, "tensorValue": self.tensor(mod.running_var)}}
            return self.module("BatchNormalization" if isinstance(mod, nn.BatchNorm1d) else "SpatialBatchNormalization", name, at,
                               weight=self.tensor(mod.weight) if mod.affine else None, bias=self.tensor(mod.bias) if mod.affine else None)
        simple = {nn.ReLU: "ReLU", nn.ReLU6: "ReLU6", nn.Tanh: "Tanh", nn.Sigmoid: "Sigmoid", nn.Softmax: "SoftMax", nn.LogSoftmax: "LogSoftMax",
                  nn.Identity: "Identity", nn.Softplus: "SoftPlus", nn.Softsign: "SoftSign"}
        for cls, short in simple.items():
            if type(mod) is cls:
                return self.module(short, name)
        if isinstance(mod, nn.Dropout):
            return self.module("Dropout", name, {"initP": self.a_double(mod.p)})
        if isinstance(mod, nn.ELU):
            return self.module("ELU", name, {"alpha": self.a_double(mod.alpha)})
        if isinstance(mod, nn.LeakyReLU):
            return self.module("LeakyReLU", name, {"negval": self.a_double(mod.negative_slope)})
        if isinstance(mod, nn.Flatten):
            return self.module("InferReshape", name, {"size": self.a_ints([-1]), "batchMode": self.a_bool(True)})
        raise NotImplementedError(f"save_bigdl: no BigDL counterpart registered for {type(mod).__name__}")


def save_bigdl(model: nn.Module, path=None, weight_path=None):
    """Serialize `model` in BigDL's module format; returns the bytes (and writes them to `path` when given).  With `weight_path`
    the definition file carries no tensor data and the weights go to that file in BigDL's separate weight-file layout
    (`saveModule(path, weightPath)`; see `bigdl_loader.read_weight_file`)."""
    import hashlib
    import struct

    from ...util import filesystem as zfs

    w = _Writer()
    root = w.convert(model.eval() if not isinstance(model, nn.Sequential) else model)
    entries = w.storage_entries
    if weight_path is not None:
        payload, kept = b"", []
        for e in entries:
            tv = e["value"]["tensorValue"]
            st = tv["storage"]
            data = np.asarray(st.get("float_data", []), dtype=">f4")
            payload += struct.pack(">iq", int(st["id"]), data.size) + data.tobytes()
            kept.append({"key": e["key"], "value": {"dataType": e["value"].get("dataType", S.DT_TENSOR),
                                                     "tensorValue": dict(tv, storage={"datatype": st.get("datatype", S.DT_FLOAT), "id": st["id"]})}})
        blob = struct.pack(">ii", 3721, len(entries)) + payload
        blob += struct.pack(">i", 16) + hashlib.md5(blob).digest()
        with zfs.open_file(weight_path, "wb") as fh:
            fh.write(blob)
        entries = kept
    root["attr"] = list(root["attr"]) + [{"key": "global_storage", "value": {"nameAttrListValue": {"name": "global_storage", "attr": entries}}}]
    buf = pw.encode(root, S.MODULE)
    if path is not None:
        with zfs.open_file(path, "wb") as fh:
            fh.write(buf)
    return buf
