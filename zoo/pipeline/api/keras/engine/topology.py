"""KerasNet / Model / Sequential: the keras-style training face.

Reference: pyzoo/zoo/pipeline/api/keras/engine/topology.py:31-403 and
zoo/src/main/scala/.../pipeline/api/keras/models/Topology.scala:65-961 -- `compile`
(string names -> optim/loss/metrics), `fit` (arrays / FeatureSet / ImageSet / TextSet),
`evaluate`, `predict`, `predict_classes`, `set_tensorboard`, `get_train_summary`,
`get_validation_summary`, `set_checkpoint`, gradient clipping, `summary`, `saveModel`.
A KerasNet here is a ``torch.nn.Module``; layers are ``zoo.pipeline.api.keras.layers``.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.nn as nn

from .....common.triggers import EveryEpoch, MaxEpoch
from .....feature.common import FeatureSet
from ..metrics import to_metric
from ..objectives import to_loss


class KerasNet(nn.Module):
    def __init__(self):
        super().__init__()
        self._engine = None
        self._compiled = None
        self._pending = {}

    # -- weights (KerasNet.get_weights / set_weights of the reference: a flat list of ndarrays) ----------------------
    def get_weights(self):
        return [p.detach().cpu().numpy() for p in self.parameters()]

    def set_weights(self, weights):
        ps = list(self.parameters())
        if len(ps) != len(weights):
            raise ValueError(f"expected {len(ps)} weight arrays, got {len(weights)}")
        with torch.no_grad():
            for p, w in zip(ps, weights):
                p.copy_(torch.as_tensor(w).reshape(p.shape))
        return self

    # -- engine plumbing -------------------------------------------------------------
    def _get_engine(self, optimizer=None):
        from ....engine import TrainingEngine

        if self._engine is None:
            self._engine = TrainingEngine(self, optimizer)
            p = self._pending
            if "tb" in p:
                self._engine.set_tensorboard(*p["tb"])
            if "ckpt" in p:
                self._engine.set_checkpoint(*p["ckpt"])
            if "clip" in p:
                self._engine.clip = p["clip"]
            if "drop" in p:
                self._engine.set_drop_module_property(*p["drop"])
        return self._engine

    def compile(self, optimizer, loss, metrics=None):
        """optimizer: name | OptimMethod | torch optimizer | dict per sub-module; loss: name | LossFunction | callable."""
        loss = to_loss(loss)
        self._compiled = dict(optimizer=optimizer, loss=loss, metrics=[to_metric(m, loss) for m in (metrics or [])])
        if self._engine is not None:  # re-compile: keep weights, rebuild optimizer state
            self._engine.close()
            self._engine = None
        self._get_engine(optimizer)
        return self

    def set_tensorboard(self, log_dir, app_name):
        self._pending["tb"] = (log_dir, app_name)
        if self._engine is not None:
            self._engine.set_tensorboard(log_dir, app_name)

    def get_train_summary(self, tag=None):
        s = self._engine.train_summary if self._engine else None
        return [] if s is None else s.read_scalar(tag)

    def get_validation_summary(self, tag=None):
        s = self._engine.val_summary if self._engine else None
        return [] if s is None else s.read_scalar(tag)

    def set_checkpoint(self, path, over_write=True):
        self._pending["ckpt"] = (path, over_write)
        if self._engine is not None:
            self._engine.set_checkpoint(path, over_write)

    def clear_gradient_clipping(self):
        self._pending.pop("clip", None)
        if self._engine is not None:
            self._engine.clear_gradient_clipping()

    def set_constant_gradient_clipping(self, min, max):
        self._pending["clip"] = ("const", float(min), float(max))
        if self._engine is not None:
            self._engine.set_constant_gradient_clipping(min, max)

    def set_gradient_clipping_by_l2_norm(self, clip_norm):
        self._pending["clip"] = ("l2", float(clip_norm))
        if self._engine is not None:
            self._engine.set_gradient_clipping_by_l2_norm(clip_norm)

    def set_drop_module_property(self, drop_percentage, max_drop_percentage, batch_size=100, warmup_iteration=200, micro_batches=4):
        """Gradient dropping / straggler mitigation of `fit` (reference: the dropPercentage family of the zoo optimizer,
        Topology.scala:1096-1100; `zoo.pipeline.straggler`)."""
        self._pending["drop"] = (drop_percentage, max_drop_percentage, batch_size, warmup_iteration, micro_batches)
        if self._engine is not None:
            self._engine.set_drop_module_property(*self._pending["drop"])
        return self

    def set_evaluate_status(self):
        self.eval()
        return self

    def regularization_loss(self):
        """Sum of the layers' weight regularizers (`W_regularizer=l2(...)`), or None when no layer has one."""
        total = None
        for m in self.modules():
            for reg, t in getattr(m, "_zoo_regularizers", ()):
                v = reg(t.float())
                total = v if total is None else total + v
        return total

    # -- fit / evaluate / predict ------------------------------------------------------
    @staticmethod
    def _as_set(x, y=None):
        if isinstance(x, FeatureSet):
            return x
        if hasattr(x, "to_feature_set"):
            return x.to_feature_set()
        if isinstance(x, (list, tuple)) and x and not isinstance(x[0], (np.ndarray, torch.Tensor)):
            return FeatureSet.sample_rdd(x)
        return FeatureSet.numpy(x, y)

    def fit(self, x, y=None, batch_size=32, nb_epoch=10, validation_split=0, validation_data=None, distributed=True):
        if self._compiled is None:
            raise RuntimeError("compile(optimizer, loss, metrics) must be called before fit")
        if validation_split and validation_data is None and isinstance(x, np.ndarray) and y is not None:
            if not 0 <= validation_split <= 1:
                raise TypeError("validation split must in range [0, 1]")
            k = int(len(x) * (1 - validation_split))
            validation_data, x, y = (x[k:], y[k:]), x[:k], y[:k]
        train = self._as_set(x, y)
        val = None
        if validation_data is not None:
            val = self._as_set(*validation_data) if isinstance(validation_data, (tuple, list)) and len(validation_data) == 2 \
                else self._as_set(validation_data)
        eng = self._get_engine(self._compiled["optimizer"])
        end = MaxEpoch(eng.state["epoch"] + nb_epoch - 1)
        trig = EveryEpoch() if (val is not None or eng.ckpt_dir is not None) else None
        eng.train(train, self._compiled["loss"], end, trig, val, self._compiled["metrics"] or (["loss"] if val is not None else None),
                  batch_size)
        return self

    def evaluate(self, x, y=None, batch_size=32):
        if self._compiled is None:
            raise RuntimeError("compile(...) must be called before evaluate")
        eng = self._get_engine(self._compiled["optimizer"])
        methods = [to_metric("loss", self._compiled["loss"])] + list(self._compiled["metrics"])
        res = eng.evaluate(self._as_set(x, y), methods, batch_size, self._compiled["loss"])
        return [r.value for r in res]

    def predict(self, x, batch_per_thread=4, distributed=True, batch_size=None):
        eng = self._get_engine((self._compiled or {}).get("optimizer"))
        bs = batch_size or max(int(batch_per_thread), 1) * 256
        return eng.predict(self._as_set(x), bs)

    def predict_classes(self, x, batch_per_thread=4, zero_based_label=True, batch_size=None):
        p = self.predict(x, batch_per_thread, batch_size=batch_size)
        c = np.asarray(p).argmax(-1)
        return c if zero_based_label else c + 1

    # -- misc ----------------------------------------------------------------------------
    def get_layer(self, name):
        for n, m in self.named_modules():
            if n == name or getattr(m, "zoo_name", None) == name:
                return m
        raise KeyError(f"no layer named {name!r}")

    @property
    def layers(self):
        return list(self.children())

    def flattened_layers(self, include_container=False):
        return [m for m in self.modules() if include_container or not list(m.children())]

    def summary(self, line_length=120, positions=(.33, .55, .67, 1.)):
        rows, total = [], 0
        for n, m in self.named_modules():
            if n and not list(m.children()):
                k = sum(p.numel() for p in m.parameters(recurse=False))
                total += k
                rows.append((f"{n} ({type(m).__name__})", m.extra_repr()[:40], k))
        w = [int(line_length * p) for p in positions]
        print("_" * line_length)
        print("Layer (type)".ljust(w[0]) + "Config".ljust(w[1] - w[0]) + "Param #")
        print("=" * line_length)
        for a, b, k in rows:
            print(a[:w[0] - 1].ljust(w[0]) + b.ljust(w[1] - w[0]) + str(k))
        print("=" * line_length)
        print(f"Total params: {total:,}")
        return total

    def save(self, path, over_write=False):
        return self.saveModel(path, None, over_write)

    def saveModel(self, modelPath, weightPath=None, over_write=False):
        if os.path.exists(modelPath) and not over_write:
            raise IOError(f"{modelPath} exists; pass over_write=True")
        torch.save(self, modelPath)

    def to_model(self):
        return self

    # ---- exporters (reference: KerasNet.saveToTf / saveToKeras2, Topology.scala:559-575 -> Net.scala NetSaver) ----------------------
    def _export_layers(self):
        layers = []
        for l in (list(self.seq) if hasattr(self, "seq") else self.layers):
            layers += l._export_layers() if isinstance(l, KerasNet) else [l]
        if not layers:
            raise ValueError("the model has no layers")
        return layers

    def save_to_tf(self, dir, python=None):
        """Frozen TensorFlow graph of this model in `dir` (`frozen_inference_graph.pb` + `graph_meta.json`, the layout TFNet reads).  The
        reference runs a generated Keras script under `python`; here the graph is written directly (`zoo.util.tf.export_tf`).  Covered:
        stacks of Dense / Activation / Dropout / Flatten layers."""
        import torch.nn as nn

        from ..layers import core as C
        from .....util.tf import export_tf

        acts = {"relu": nn.ReLU, "sigmoid": nn.Sigmoid, "tanh": nn.Tanh, "softmax": lambda: nn.Softmax(-1), "log_softmax": lambda: nn.LogSoftmax(-1),
                "softplus": nn.Softplus, "elu": nn.ELU, "relu6": nn.ReLU6}
        mods = []
        for l in self._export_layers():
            if isinstance(l, C.Dense):
                mods.append(l.dense)
                if l.activation not in (None, "linear", "relu"):
                    mods.append(acts[l.activation]())
            elif isinstance(l, C.Activation):
                if l.activation not in (None, "linear"):
                    mods.append(acts[l.activation]())
            elif isinstance(l, C.Dropout):
                mods.append(nn.Dropout(l.p))
            elif isinstance(l, C.Flatten):
                mods.append(nn.Flatten())
            else:
                raise NotImplementedError(f"save_to_tf: no TensorFlow graph is written for layer {type(l).__name__} "
                                          "(covered: Dense, Activation, Dropout, Flatten)")
        shape = self.get_input_shape()
        example = torch.zeros((1,) + tuple(int(d) for d in shape[1:]))
        return export_tf(nn.Sequential(*mods), dir, example_input=example)

    saveToTf = save_to_tf

    def save_to_keras2(self, file_path, python=None):
        """What the reference produces with a Keras installation (a script that builds the Keras-2 model, sets the weights and calls
        `model.save`): here the two inputs of that step are written -- `<file_path>` = the weights as a Keras-layout HDF5 file
        (`layer_names` / `weight_names`, Dense kernels as [in, out]; `zoo.util.hdf5`) and `<file_path>.py` = the Keras-2 script that
        rebuilds the architecture and loads them.  With `python` given (an interpreter that has Keras) the script is run.  Covered:
        Dense, Activation, Dropout, Flatten, Reshape, Permute, LSTM, GRU, Embedding."""
        import subprocess

        import numpy as np

        from .....util.hdf5 import write_hdf5

        lines = ["from tensorflow.keras.models import Sequential", "from tensorflow.keras.layers import *", "",
                 "model = Sequential(name=%r)" % (getattr(self, "zoo_name", None) or "model")]
        tree, names = {}, []
        first = True
        for k, l in enumerate(self._export_layers()):
            cname = type(l).__name__
            lname = getattr(l, "zoo_name", None)
            lname = lname if isinstance(lname, str) and lname else f"{cname.lower()}_{k}"
            extra = ", input_shape=%r" % (tuple(int(d) for d in self.get_input_shape()[1:]),) if first else ""
            first = False
            if cname == "Dense":
                code = f"Dense({l.output_dim}, activation={l.activation!r}, use_bias={bool(l.use_bias)}, name={lname!r}{extra})"
            elif cname == "Activation":
                code = f"Activation({l.activation!r}, name={lname!r}{extra})"
            elif cname == "Dropout":
                code = f"Dropout({l.p}, name={lname!r}{extra})"
            elif cname == "Flatten":
                code = f"Flatten(name={lname!r}{extra})"
            elif cname == "Reshape":
                code = f"Reshape({tuple(l.target_shape)!r}, name={lname!r}{extra})"
            elif cname == "Permute":
                code = f"Permute({tuple(l.dims)!r}, name={lname!r}{extra})"
            elif cname in ("LSTM", "GRU"):
                code = (f"{cname}({l.output_dim}, activation={getattr(l, 'activation', 'tanh')!r}, recurrent_activation={l.inner_activation!r}, "
                        f"return_sequences={bool(l.return_sequences)}, go_backwards={bool(l.go_backwards)}, name={lname!r}{extra})")
            elif cname == "Embedding":
                code = f"Embedding({l.input_dim}, {l.output_dim}, name={lname!r}{extra})"
            else:
                raise NotImplementedError(f"save_to_keras2: no Keras-2 constructor is written for layer {cname}")
            lines.append(f"model.add({code})")
            params = [(n, p.detach().float().cpu().numpy()) for n, p in l.named_parameters()]
            if params:
                arrays = {}
                for n, a in params:
                    a = a.T.copy() if cname == "Dense" and a.ndim == 2 else a          # Keras kernels are [in, out]
                    arrays[f"{n.replace('.', '_')}:0"] = a
                tree[lname] = {"__attrs__": {"weight_names": np.array([f"{lname}/{n}".encode() for n in arrays])}, lname: arrays}
                names.append(lname)
        write_hdf5(file_path, tree, attrs={"layer_names": np.array([n.encode() for n in names]), "backend": "tensorflow", "keras_version": "2"})
        full = str(file_path)[:-3] + "_full.h5" if str(file_path).endswith(".h5") else str(file_path) + "_full.h5"
        lines += ["", f"model.load_weights({str(file_path)!r}, by_name=True)", f"model.save({full!r})", ""]
        script = str(file_path) + ".py"
        with open(script, "w") as f:
            f.write("\n".join(lines))
        if python:
            subprocess.run([python, script], check=True)
        return file_path, script

    saveToKeras2 = save_to_keras2


class Sequential(KerasNet):
    """Linear stack of layers: `model.add(layer)`.  The first layer needs `input_shape=`; every later
    layer is built from the previous layer's output shape (Topology.scala:827-961)."""

    def __init__(self, layers=None, name=None):
        super().__init__()
        self.seq = nn.ModuleList()
        self.zoo_name = name
        self._out_shape = None
        self._shape_known = True
        for l in layers or []:
            self.add(l)

    def add(self, layer):
        from ..layers.base import ZooKerasLayer

        if isinstance(layer, ZooKerasLayer):
            shp = self._out_shape if len(self.seq) else layer._declared_input_shape
            if len(self.seq) == 0 and shp is None and not layer.built:
                raise ValueError(f"{layer.zoo_name}: the first layer of a Sequential model needs input_shape=")
            if self._shape_known and shp is not None:
                self._out_shape = layer.ensure_built(shp)
            elif layer.built:
                self._out_shape, self._shape_known = layer.output_shape_, layer.output_shape_ is not None
        elif isinstance(layer, Sequential) and layer._shape_known:
            self._out_shape = layer._out_shape
        else:
            self._out_shape, self._shape_known = None, False      # plain torch module: shapes resolve at the first call
        self.seq.append(layer)
        return self

    def get_output_shape(self):
        return None if self._out_shape is None else ((None,) + tuple(self._out_shape) if not isinstance(self._out_shape, list)
                                                     else [(None,) + tuple(s) for s in self._out_shape])

    def get_input_shape(self):
        first = self.seq[0] if len(self.seq) else None
        return first.get_input_shape() if hasattr(first, "get_input_shape") else None

    # a Sequential can itself be used as a layer of another model / called on Variables
    _declared_input_shape = None
    built = True

    @property
    def output_shape_(self):
        return self._out_shape

    def ensure_built(self, input_shape):
        return self._out_shape

    def __call__(self, *args, **kwargs):
        from ...autograd import Variable

        flat = args[0] if (len(args) == 1 and isinstance(args[0], (list, tuple))) else args
        if flat and all(isinstance(a, Variable) for a in flat):
            return Variable.from_layer(self, list(flat))
        return super().__call__(*args, **kwargs)

    def forward(self, *xs):
        x = xs[0] if len(xs) == 1 else list(xs)
        for l in self.seq:
            x = l(*x) if isinstance(x, (list, tuple)) and getattr(l, "multi_input", False) else l(x)
        return x

    def logits_kind(self):
        last = self.seq[-1] if len(self.seq) else None
        return getattr(last, "final_activation_kind", None) if hasattr(last, "strip_final_activation") else None

    def forward_logits(self, *xs):
        """Everything but a trailing softmax/sigmoid activation (enables the fused loss kernels)."""
        last = self.seq[-1] if len(self.seq) else None
        strip = getattr(last, "strip_final_activation", None)
        x = xs[0] if len(xs) == 1 else list(xs)
        for l in list(self.seq)[:-1]:
            x = l(*x) if isinstance(x, (list, tuple)) and getattr(l, "multi_input", False) else l(x)
        if last is None:
            return x
        if not getattr(last, "built", True):
            return last(x)
        return strip(x) if strip is not None else last(x)


class Model(KerasNet):
    """Functional graph container: `Model(input, output)` over symbolic `Variable`s
    (zoo.pipeline.api.autograd), or `Model(fn=callable)`."""

    def __init__(self, input=None, output=None, name=None, fn=None):
        super().__init__()
        self.zoo_name = name
        if fn is not None:
            self.fn = fn
            self.graph = None
        else:
            from ...autograd import GraphModule

            self.graph = GraphModule(input, output)
            self.fn = None

    def forward(self, *xs):
        if self.graph is not None:
            return self.graph(*xs)
        return self.fn(*xs)

    def get_input_shape(self):
        s = [v.get_output_shape() for v in self.graph.inputs]
        return s[0] if len(s) == 1 else s

    def get_output_shape(self):
        s = [v.get_output_shape() for v in self.graph.outputs]
        return s[0] if self.graph.single_output else s

    def __call__(self, *args, **kwargs):
        from ...autograd import Variable

        flat = args[0] if (len(args) == 1 and isinstance(args[0], (list, tuple))) else args
        if flat and all(isinstance(a, Variable) for a in flat):
            return Variable.from_layer(self, list(flat))
        return super().__call__(*args, **kwargs)

    def new_graph(self, outputs):
        """Sub-graph ending at the named layers (Net.scala `newGraph`)."""
        from ...autograd import GraphModule

        names = [outputs] if isinstance(outputs, str) else list(outputs)
        sel = []
        for n in names:
            vs = [v for v in self.graph.order if v.module is not None and getattr(v.module, "zoo_name", None) == n]
            if not vs:
                raise KeyError(f"no layer named {n!r}")
            sel.append(vs[-1])
        m = Model.__new__(Model)
        KerasNet.__init__(m)
        m.zoo_name, m.fn = None, None
        m.graph = GraphModule(self.graph.inputs, sel[0] if len(sel) == 1 else sel)
        return m

    def freeze_up_to(self, names):
        """Stop training every layer up to (and including) the named ones (Net.scala `freezeUpTo`)."""
        names = [names] if isinstance(names, str) else list(names)
        sub = self.new_graph(names)
        for p in sub.parameters():
            p.requires_grad_(False)
        return self

    def unfreeze(self):
        for p in self.parameters():
            p.requires_grad_(True)
        return self


# names the reference defines in this module (engine/topology.py: Input, InputLayer, Merge, merge)
from ..layers.base import Input, InputLayer  # noqa: E402,F401
from ..layers.core import Merge, merge  # noqa: E402,F401
