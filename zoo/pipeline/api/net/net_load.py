"""Net.load* (reference: pyzoo/zoo/pipeline/api/net/net_load.py:69-140, zoo/.../pipeline/api/Net.scala:51-446).

`Net.load` / `load_bigdl` read a model written by this framework (`ZooModel.save_model`, `KerasNet.saveModel`);
`load_torch` reads a TorchScript archive or a pickled module.  The foreign formats of the reference (BigDL
protobuf, Torch7 `.t7`, Caffe prototxt/caffemodel, Keras-1 json/hdf5, TensorFlow frozen graphs) need engines
are imported where a file format can be read without its engine: Caffe (zoo.models.caffe), TensorFlow frozen graphs
(zoo.tfpark.graphdef), Keras-1 JSON definitions; a BigDL protobuf or a Torch7 `.t7` file raises a `NotImplementedError` that names
the supported route."""
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

        with open(model_path, "rb") as f:
            head = f.read(2)
        if head not in (b"PK", b"\x80\x02", b"\x80\x03", b"\x80\x04", b"\x80\x05"):      # not a torch archive: BigDL's own format
            return Net.load_bigdl(model_path, weight_path)
        try:
            m = ZooModel.load_model(model_path, weight_path)
        except Exception:
            m = torch.load(model_path, map_location="cpu", weights_only=False)
            if weight_path is not None:
                m.load_state_dict(torch.load(weight_path, map_location="cpu"))
        return GraphNet(m)

    @staticmethod
    def load_bigdl(model_path, weight_path=None, bigdl_type="float"):
        """A model saved by BigDL / Analytics Zoo (`saveModule` / `saveModel`: BigDL's protobuf format, decoded by
        `zoo.models.bigdl`) or by this framework (`ZooModel.save_model`: a torch archive)."""
        with open(model_path, "rb") as f:
            head = f.read(2)
        if head not in (b"PK", b"\x80\x02", b"\x80\x03", b"\x80\x04", b"\x80\x05"):
            from ....models.bigdl import load_bigdl as _load

            return GraphNet(_load(model_path, weight_path))
        return Net.load(model_path, weight_path)

    @staticmethod
    def save_bigdl(model, path, weight_path=None):
        """Write a feed-forward torch model in BigDL's module format (loadable by BigDL / Analytics Zoo and by `load_bigdl`);
        with `weight_path` the weights go to BigDL's separate weight file (`saveModule(path, weightPath)`)."""
        from ....models.bigdl import save_bigdl as _save

        _save(getattr(model, "model", model) if not isinstance(model, torch.nn.Module) else model, path, weight_path)
        return path

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

        return GraphNet(_load(def_path, model_path)).eval()     # inference mode, as a freshly loaded Caffe net; training code calls .train()

    @staticmethod
    def load_keras(json_path=None, hdf5_path=None, by_name=False):
        """A Keras-1.2.2 model definition (`model.to_json()`) -> zoo Keras-style Sequential (the zoo layers keep Keras-1 argument
        names, so the layer configs map one to one); weights from `hdf5_path` (a Keras `save_weights` HDF5 file; read with h5py, or with
        `zoo.util.hdf5` when h5py is absent), or from an `.npz` file holding
        the arrays of `model.get_weights()` in order (reference: Net.load_keras, net_load.py:117-140)."""
        import inspect
        import json

        import numpy as np

        from ..keras import layers as L
        from ..keras.engine.topology import Sequential

        if json_path is None:
            raise NotImplementedError("Net.load_keras needs the model definition (json_path); an hdf5 file alone carries a pickled "
                                      "Keras object graph that only Keras can rebuild")
        spec = json.load(open(json_path))
        if spec.get("class_name") != "Sequential":
            raise NotImplementedError(f"Net.load_keras: only Sequential definitions are imported (got {spec.get('class_name')!r}); build "
                                      "functional graphs with zoo.pipeline.api.keras.models.Model")
        model = Sequential()
        for entry in spec["config"]:
            cname, cfg = entry["class_name"], dict(entry.get("config", {}))
            cls = getattr(L, cname, None)
            if cls is None:
                raise NotImplementedError(f"Net.load_keras: no zoo layer named {cname}")
            if cfg.get("batch_input_shape"):
                cfg["input_shape"] = tuple(cfg.pop("batch_input_shape")[1:])
            for k in ("W_regularizer", "b_regularizer", "activity_regularizer", "W_constraint", "b_constraint"):
                cfg.pop(k, None)          # Keras serialises these as dicts / None; the zoo layers take objects
            accepted = inspect.signature(cls.__init__).parameters
            kwargs = {k: (tuple(v) if isinstance(v, list) else v) for k, v in cfg.items() if k in accepted and v is not None}
            model.add(cls(**kwargs))
        if hdf5_path is not None:
            if str(hdf5_path).endswith(".npz"):
                arrays = [v for _, v in sorted(np.load(hdf5_path).items(), key=lambda kv: int("".join(ch for ch in kv[0] if ch.isdigit()) or 0))]
            else:
                try:
                    import h5py as h5
                except ImportError:
                    from ....util import hdf5 as h5        # the built-in reader of the classic HDF5 layout (what Keras / h5py write)
                arrays = []
                with h5.File(hdf5_path, "r") as f:
                    g = f["model_weights"] if "model_weights" in f else f
                    for lname in [n.decode() if isinstance(n, bytes) else n for n in g.attrs["layer_names"]]:
                        for wname in [n.decode() if isinstance(n, bytes) else n for n in g[lname].attrs["weight_names"]]:
                            arrays.append(np.asarray(g[lname][wname]))
            params = [p for p in model.parameters()]
            if len(params) != len(arrays):
                raise ValueError(f"Net.load_keras: the definition has {len(params)} weight tensors, the file {len(arrays)}")
            with torch.no_grad():
                for p, a in zip(params, arrays):
                    t = torch.from_numpy(np.asarray(a, dtype=np.float32))
                    if t.shape != p.shape and t.t().shape == p.shape:
                        t = t.t()             # Keras Dense kernels are [in, out]
                    p.copy_(t.reshape(p.shape))
        return model

    @staticmethod
    def load_tf(path, inputs=None, outputs=None, byte_order="little_endian", bin_file=None, bigdl_type="float"):
        """A frozen GraphDef (`.pb` + input / output names) or an export folder (frozen_inference_graph.pb + graph_meta.json) ->
        GraphNet over the imported torch graph (zoo.tfpark.graphdef; Net.scala loadTF)."""
        import os

        from .tfnet import TFNet

        if os.path.isdir(path) and os.path.exists(os.path.join(path, "saved_model.pb")):      # a SavedModel directory
            return GraphNet(TFNet.from_saved_model(path, inputs=inputs, outputs=outputs))
        return GraphNet(TFNet(path, inputs, outputs))


class JavaToPython:
    """py4j class-name mapping of the reference (net_load.py:30-66); there is no JVM here, kept as a name."""

    def __init__(self, jvalue=None, bigdl_type="float"):
        self.jvalue = jvalue

    def get_python_class(self):
        return type(self.jvalue)
