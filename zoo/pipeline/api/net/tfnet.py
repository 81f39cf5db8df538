"""TFNet: a frozen TensorFlow graph as a layer / predictor (reference: pyzoo/zoo/tfpark/tfnet.py:52-302 and
zoo/.../pipeline/api/net/TFNet.scala:56-716, which run the graph through libtensorflow JNI).

There is no TensorFlow runtime on this stack: the frozen GraphDef is imported into a torch module
(:mod:`zoo.tfpark.graphdef`), so a TFNet runs on the GPU like any other model, can sit inside a Sequential, and serves
through `InferenceModel`.  Supported sources are files: a frozen `.pb` plus input / output names, or the reference's export
folder (`frozen_inference_graph.pb` + `graph_meta.json`, written by `zoo.util.tf.export_tf`).  Constructors that need a live
`tf.Session` (`from_session`) raise: there is no session to read variables from.
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch
import torch.nn as nn


class TFNet(nn.Module):
    def __init__(self, path, input_names=None, output_names=None, tf_session_config=None, bigdl_type="float"):
        super().__init__()
        from ....tfpark.graphdef import GraphDefModule, read_graphdef

        trainable = None
        if os.path.isdir(path):
            meta = json.load(open(os.path.join(path, "graph_meta.json")))
            input_names = input_names or meta["input_names"]
            output_names = output_names or meta["output_names"]
            trainable = meta.get("variables")        # a training export (export_tf(generate_backward=True)): its variables are trainable here too
            path = os.path.join(path, "frozen_inference_graph.pb")
        self.path = path
        self.graph = GraphDefModule(read_graphdef(path), input_names, output_names, trainable=trainable)
        self.input_names, self.output_names = self.graph.input_names, self.graph.output_names

    @staticmethod
    def from_export_folder(folder, tf_session_config=None):
        """Folder written by `export_tf` (frozen_inference_graph.pb + graph_meta.json)."""
        if not os.path.isdir(folder):
            raise ValueError(f"{folder} does not exist")
        return TFNet(folder, tf_session_config=tf_session_config)

    @staticmethod
    def from_frozen_graph(path, input_names, output_names):
        return TFNet(path, input_names, output_names)

    @staticmethod
    def from_saved_model(model_path, tag=None, signature=None, inputs=None, outputs=None, tf_session_config=None, init_op=None):
        """A SavedModel directory (`saved_model.pb` + `variables/`): the meta graph tagged `tag` ("serve"), variables restored from
        the bundle and frozen into the graph, run between `inputs` / `outputs` or the tensors of `signature` ("serving_default").
        Reference: TFNet.fromSavedModel, TFNet.scala:612-716.  A folder written by `export_tf` is accepted too."""
        if os.path.exists(os.path.join(model_path, "frozen_inference_graph.pb")) and not os.path.exists(os.path.join(model_path, "saved_model.pb")):
            return TFNet(model_path, inputs, outputs)
        from ....tfpark.graphdef import GraphDefModule
        from ....tfpark.saved_model import load_saved_model

        nodes, ins, outs = load_saved_model(model_path, tag, signature, inputs, outputs)
        net = TFNet.__new__(TFNet)
        nn.Module.__init__(net)
        net.path = model_path
        net.graph = GraphDefModule(nodes, ins, outs)
        net.input_names, net.output_names = net.graph.input_names, net.graph.output_names
        return net

    @staticmethod
    def from_session(sess, inputs, outputs, generate_backward=False, allow_non_differentiable_input=True):
        raise NotImplementedError("TFNet.from_session reads variables from a live tf.Session; there is no TensorFlow runtime "
                                  "here -- export the graph with export_tf in a TF environment and use TFNet.from_export_folder")

    def forward(self, *xs):
        return self.graph(*xs)

    @torch.no_grad()
    def predict(self, x, batch_per_thread=32, distributed=False):
        """ndarray / tensor (or a list of them for multi-input graphs) -> ndarray of the first output, batched."""
        xs = x if isinstance(x, (list, tuple)) else [x]
        xs = [torch.as_tensor(np.asarray(t)) if not torch.is_tensor(t) else t for t in xs]
        dev = next(self.graph.buffers(), torch.zeros(1)).device
        outs = []
        for i in range(0, xs[0].shape[0], max(1, batch_per_thread)):
            o = self.graph(*[t[i:i + batch_per_thread].to(dev) for t in xs])
            outs.append(o if torch.is_tensor(o) else o[0])
        return torch.cat(outs).float().cpu().numpy()


class TFNetForInference(TFNet):
    """A SavedModel loaded for inference with its variables restored (reference: TFNetForInference.scala -- the variant of TFNet
    that keeps variables in the graph and assigns them from the checkpoint instead of freezing them).  Here restoring and freezing
    are the same step (`zoo.tfpark.saved_model`), so this is `TFNet.from_saved_model` under the reference's class name."""

    def __new__(cls, model_path, tag=None, signature=None, inputs=None, outputs=None, **_ignored):
        net = TFNet.from_saved_model(model_path, tag=tag, signature=signature, inputs=inputs, outputs=outputs)
        net.__class__ = cls
        return net

    def __init__(self, *a, **kw):       # built in __new__
        pass

    @staticmethod
    def from_saved_model(model_path, tag=None, signature=None, inputs=None, outputs=None, **kw):
        return TFNetForInference(model_path, tag, signature, inputs, outputs)

    fromSavedModel = from_saved_model


class GraphRunner:
    """Runs named tensors of a GraphDef (reference: tfpark/GraphRunner.scala: `run(input, inputNames, outputNames, targets)` over a
    TF session).  One imported module is kept per (inputs, outputs) pair."""

    def __init__(self, graph_def, restore_op=None, restore_path_placeholder=None, save_op=None, save_path_placeholder=None, config=None):
        from ....tfpark.graphdef import read_graphdef

        self.nodes = read_graphdef(graph_def)
        self._modules = {}

    def run(self, inputs, input_names, output_names, targets=None):
        from ....tfpark.graphdef import GraphDefModule

        key = (tuple(input_names), tuple(output_names))
        if key not in self._modules:
            self._modules[key] = GraphDefModule(self.nodes, list(input_names), list(output_names))
        out = self._modules[key](*inputs)
        return list(out) if isinstance(out, tuple) else [out]
