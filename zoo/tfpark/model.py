"""tfpark.KerasModel (reference: pyzoo/zoo/tfpark/model.py:34-449): `fit / evaluate / predict / save_weights /
load_weights / save_model` with `distributed=True|False`, over a compiled model.  The reference wraps a tf.keras model;
here the wrapped model is a zoo `KerasNet` (or any torch module + compile arguments), trained by the engine."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from ..pipeline.api.keras.engine.topology import KerasNet
from .tf_dataset import TFDataset


class _Wrapped(KerasNet):
    def __init__(self, module):
        super().__init__()
        self.module = module

    def forward(self, *xs):
        return self.module(*xs)


class KerasModel:
    def __init__(self, model, model_dir=None, optimizer=None, loss=None, metrics=None):
        if isinstance(model, KerasNet):
            self.model = model
        elif isinstance(model, nn.Module):
            self.model = _Wrapped(model)
        else:
            raise TypeError("tfpark.KerasModel wraps a zoo KerasNet or a torch module (tf.keras models need TensorFlow, which is not part of zoo-b200)")
        if self.model._compiled is None:
            if optimizer is None or loss is None:
                raise ValueError("pass a compiled KerasNet, or optimizer= and loss=")
            self.model.compile(optimizer, loss, metrics)
        self.model_dir = model_dir

    @property
    def metrics_names(self):
        return ["loss"] + [m.name for m in self.model._compiled["metrics"]]

    def get_weights(self):
        return [p.detach().cpu().numpy() for p in self.model.parameters()]

    def set_weights(self, weights):
        with torch.no_grad():
            for p, w in zip(self.model.parameters(), weights):
                p.copy_(torch.as_tensor(w).reshape(p.shape))

    def save_weights(self, filepath, overwrite=True, save_format=None):
        """`.h5` / `.hdf5` (or save_format="h5"): a Keras-layout HDF5 file -- root attribute `layer_names`, one group per layer with
        `weight_names` and its arrays (written by `zoo.util.hdf5`); anything else: a torch state_dict."""
        import os

        if os.path.exists(filepath) and not overwrite:
            raise FileExistsError(filepath)
        if save_format in ("h5", "hdf5") or str(filepath).endswith((".h5", ".hdf5")):
            import numpy as np

            from ..util.hdf5 import write_hdf5

            tree, layers = {}, []
            for lname, mod in self.model.named_modules():
                params = [(n, p) for n, p in mod.named_parameters(recurse=False)] + [(n, b) for n,
                        b in mod.named_buffers(recurse=False) if b.is_floating_point()]
                if not params:
                    continue
                key = lname.replace(".", "_") or "model"
                names = [f"{key}/{n}:0" for n, _ in params]
                tree[key] = {"__attrs__": {"weight_names": np.array([n.encode() for n in names])},
                        key: {f"{n}:0": p.detach().float().cpu().numpy() for n, p in params}}
                layers.append(key)
            write_hdf5(filepath, tree, attrs={"layer_names": np.array([n.encode() for n in layers]), "backend": "torch", "keras_version": "zoo-b200"})
            return
        torch.save(self.model.state_dict(), filepath)

    def load_weights(self, filepath, by_name=False):
        with open(filepath, "rb") as f:
            head = f.read(2048)
        if b"\x89HDF\r\n\x1a\n" in head:
            import numpy as np

            try:
                import h5py as h5
            except ImportError:
                from ..util import hdf5 as h5
            arrays = {}
            with h5.File(filepath, "r") as f:
                g = f["model_weights"] if "model_weights" in f else f
                for lname in [n.decode() if isinstance(n, bytes) else n for n in g.attrs["layer_names"]]:
                    for wname in [n.decode() if isinstance(n, bytes) else n for n in g[lname].attrs["weight_names"]]:
                        arrays[wname] = np.asarray(g[lname][wname])
            own = {}
            for lname, mod in self.model.named_modules():
                key = lname.replace(".", "_") or "model"
                for n, t in list(mod.named_parameters(recurse=False)) + list(mod.named_buffers(recurse=False)):
                    own[f"{key}/{n}:0"] = t
            if by_name or set(arrays) <= set(own):
                pairs = [(own[k], v) for k, v in arrays.items() if k in own]
            else:                                           # a file written by Keras itself: assign in order
                targets = [p for p in self.model.parameters()]
                if len(targets) != len(arrays):
                    raise ValueError(f"the model has {len(targets)} weight tensors, the file {len(arrays)}")
                pairs = list(zip(targets, arrays.values()))
            with torch.no_grad():
                for t, a in pairs:
                    v = torch.from_numpy(np.asarray(a, dtype=np.float32))
                    t.copy_((v.t() if v.shape != t.shape and v.dim() == 2 and v.t().shape == t.shape else v).reshape(t.shape))
            return
        self.model.load_state_dict(torch.load(filepath, map_location="cpu"))

    def save_model(self, path):
        torch.save(self.model, path)

    @staticmethod
    def load_model(path):
        m = torch.load(path, map_location="cpu", weights_only=False)
        km = KerasModel.__new__(KerasModel)
        km.model, km.model_dir = m, None
        return km

    @staticmethod
    def _unpack(x, y):
        if isinstance(x, TFDataset):
            return x.feature_set, None, (x.batch_size if x.batch_size > 0 else x.batch_per_thread), x.val_set
        return x, y, None, None

    def fit(self, x=None, y=None, batch_size=None, epochs=1, validation_split=0.0, validation_data=None, distributed=False, **kw):
        data, y, bs, val = self._unpack(x, y)
        self.model.fit(data, y, batch_size=batch_size or bs or 32, nb_epoch=epochs, validation_split=validation_split,
                       validation_data=validation_data if validation_data is not None else val)
        return self

    def evaluate(self, x=None, y=None, batch_per_thread=None, distributed=False, batch_size=None):
        data, y, bs, _ = self._unpack(x, y)
        return self.model.evaluate(data, y, batch_size=batch_size or bs or 32)

    def predict(self, x, batch_per_thread=None, distributed=False, batch_size=None):
        data, _, bs, _ = self._unpack(x, None)
        return np.asarray(self.model.predict(data, batch_size=batch_size or max(bs or 1, 1) * 64))

    def train_on_batch(self, x, y=None, sample_weight=None, class_weight=None, reset_metrics=True):
        self.fit(x, y, batch_size=len(x), epochs=1)
        return self.evaluate(x, y, batch_size=len(x))

    def test_on_batch(self, x, y=None, sample_weight=None, reset_metrics=True):
        return self.evaluate(x, y, batch_size=len(x))

    def predict_on_batch(self, x):
        return self.predict(x, batch_size=len(x))
