"""ZooModel / KerasZooModel: base classes of the built-in models.

Reference: pyzoo/zoo/models/common/zoo_model.py:32-130 and
zoo/src/main/scala/.../models/common/ZooModel.scala:38-154 (`saveModel(path, weightPath)`,
`loadModel`, `predictClasses`, `summary`) plus KerasZooModel.scala:32-183 (the
compile/fit/evaluate/predict/setTensorBoard/setCheckpoint/clipping passthroughs to KerasNet).
Here a model *is* a ``torch.nn.Module``; the KerasNet behaviour comes from
:class:`zoo.pipeline.api.keras.engine.topology.KerasNet`, which owns the trainer.
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch

from ...pipeline.api.keras.engine.topology import KerasNet


class ZooModel(KerasNet):
    """Base of every built-in model: constructor-argument capture + save/load."""

    _registry: dict = {}

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        ZooModel._registry[cls.__name__] = cls

    def _record_config(self, **cfg):
        self._zoo_config = {k: (list(v) if isinstance(v, tuple) else v) for k, v in cfg.items()}

    def get_config(self):
        return dict(getattr(self, "_zoo_config", {}))

    # -- persistence (ZooModel.scala:78-83,149-153) ---------------------------------
    def save_model(self, path, weight_path=None, over_write=False):
        """Write `{class, config}` + weights.  With `weight_path` the definition and the weights go
        to two files like BigDL's `saveModule(path, weightPath)`; otherwise one file holds both."""
        if os.path.exists(path) and not over_write:
            raise IOError(f"{path} already exists; pass over_write=True")
        os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
        state = {k: v.detach().cpu() for k, v in self.state_dict().items()}
        meta = {"format": "zoo-b200/1", "class": type(self).__name__, "config": self.get_config()}
        if weight_path is None:
            torch.save({"meta": meta, "state_dict": state}, path)
        else:
            if os.path.exists(weight_path) and not over_write:
                raise IOError(f"{weight_path} already exists; pass over_write=True")
            torch.save({"meta": meta, "state_dict": None}, path)
            torch.save(state, weight_path)

    saveModel = save_model

    @classmethod
    def load_model(cls, path, weight_path=None, **kw):
        blob = torch.load(path, map_location="cpu", weights_only=False)
        meta = blob["meta"]
        klass = ZooModel._registry.get(meta["class"], cls)
        model = klass(**meta["config"])
        state = blob["state_dict"] if weight_path is None else torch.load(weight_path, map_location="cpu", weights_only=False)
        if state is not None:
            model.load_state_dict(state)
        return model

    loadModel = load_model

    # -- inference helpers ----------------------------------------------------------
    def predict_classes(self, x, batch_size=32, zero_based_label=True):
        """Arg-max class of `predict(x)`; labels start at 0 (or 1 with zero_based_label=False)."""
        probs = self.predict(x, batch_size=batch_size)
        cls_ = np.asarray(probs).argmax(axis=-1)
        return cls_ if zero_based_label else cls_ + 1

    def set_evaluate_status(self):
        self.eval()
        return self


class KerasZooModel(ZooModel):
    """Keras-style built-in model (all KerasNet methods are inherited)."""


ZooModelCreator = ZooModel
