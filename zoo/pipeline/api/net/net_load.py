"""Net.load* (reference: pyzoo/zoo/pipeline/api/net/net_load.py:69-140, zoo/.../pipeline/api/Net.scala:51-446).

`Net.load` / `load_bigdl` read a model written by this framework (`ZooModel.save_model`, `KerasNet.saveModel`);
`load_torch` reads a TorchScript archive or a pickled module.  The foreign formats of the reference (BigDL
protobuf, Torch7 `.t7`, Caffe prototxt/caffemodel, Keras-1 json/hdf5, TensorFlow frozen graphs) need engines
that are not on this box; those loaders raise a `NotImplementedError` that names the supported route."""
from __future__ import annotations

import torch

from .graph_net import GraphNet


def _unsupported(what):
    raise NotImplementedError(f"{what}: this format needs an engine that is not part of zoo-b200 (no BigDL/Caffe/TensorFlow/Keras-1 "
                              "runtime on the box); convert the model to TorchScript and use Net.load_torch(path)")


class Net:
    @staticmethod
    def load(model_path, weight_path=None, bigdl_type="float"):
        from ....models.common.zoo_model import ZooModel

        try:
            m = ZooModel.load_model(model_path, weight_path)
        except Exception:
            m = torch.load(model_path, map_location="cpu", weights_only=False)
            if weight_path is not None:
                m.load_state_dict(torch.load(weight_path, map_location="cpu"))
        return GraphNet(m)

    @staticmethod
    def load_bigdl(model_path, weight_path=None, bigdl_type="float"):
        with open(model_path, "rb") as f:
            head = f.read(2)
        if head not in (b"PK", b"\x80\x02", b"\x80\x03", b"\x80\x04", b"\x80\x05"):
            _unsupported("Net.load_bigdl on a BigDL protobuf file")
        return Net.load(model_path, weight_path)

    @staticmethod
    def load_torch(path, bigdl_type="float"):
        from .torch_net import TorchNet

        try:
            return TorchNet(torch.jit.load(path, map_location="cpu"))
        except Exception:
            return TorchNet(torch.load(path, map_location="cpu", weights_only=False))

    @staticmethod
    def load_caffe(def_path, model_path, bigdl_type="float"):
        """prototxt + caffemodel -> GraphNet over the converted torch graph (zoo.models.caffe; Net.scala loadCaffe)."""
        from ....models.caffe import load_caffe as _load

        return GraphNet(_load(def_path, model_path))

    @staticmethod
    def load_keras(json_path=None, hdf5_path=None, by_name=False):
        _unsupported("Net.load_keras")

    @staticmethod
    def load_tf(path, inputs=None, outputs=None, byte_order="little_endian", bin_file=None, bigdl_type="float"):
        _unsupported("Net.load_tf")


class JavaToPython:
    """py4j class-name mapping of the reference (net_load.py:30-66); there is no JVM here, kept as a name."""

    def __init__(self, jvalue=None, bigdl_type="float"):
        self.jvalue = jvalue

    def get_python_class(self):
        return type(self.jvalue)
