"""Freezing a graph (reference examples: pyzoo/zoo/examples/tensorflow/freeze_checkpoint/freeze_checkpoint.py and
freeze_saved_model/freeze.py: restore variables, `convert_variables_to_constants`, `remove_training_nodes`, write the .pb TFNet
loads).  The same three steps on a GraphDef that still has variable nodes, using `zoo.util.tf_graph_util` (no TensorFlow)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import _common  # noqa: F401
import numpy as np
import torch

from zoo.pipeline.api.net import TFNet
from zoo.tfpark.graphdef import read_graphdef, write_graphdef
from zoo.util import tf_graph_util as gu

f32 = np.float32
training_graph = write_graphdef(None, [
    ("x", "Placeholder", [], {"dtype": f32}),
    ("dense/kernel", "VariableV2", [], {"dtype": f32}), ("dense/kernel/read", "Identity", ["dense/kernel"], {"T": f32}),
    ("dense/bias", "VariableV2", [], {"dtype": f32}), ("dense/bias/read", "Identity", ["dense/bias"], {"T": f32}),
    ("dense/MatMul", "MatMul", ["x", "dense/kernel/read"], {"T": f32, "transpose_a": False, "transpose_b": False}),
    ("dense/BiasAdd", "BiasAdd", ["dense/MatMul", "dense/bias/read"], {"T": f32, "data_format": "NHWC"}),
    ("check", "CheckNumerics", ["dense/BiasAdd"], {"T": f32}),
    ("probs", "Softmax", ["check"], {"T": f32}),
    ("unused/branch", "Relu", ["dense/MatMul"], {"T": f32}),
])
rng = np.random.RandomState(0)
checkpoint = {"dense/kernel": rng.randn(5, 3).astype(f32), "dense/bias": rng.randn(3).astype(f32)}     # what the checkpoint holds

frozen = gu.convert_variables_to_constants(checkpoint, training_graph, ["probs"])
slim = gu.remove_training_nodes(frozen, protected_nodes=["probs"])
path = os.path.join(tempfile.mkdtemp(), "frozen.pb")
with open(path, "wb") as fh:
    fh.write(slim)
print("nodes: %d in the training graph -> %d frozen -> %d after removing training nodes" %
      (len(read_graphdef(training_graph)), len(read_graphdef(frozen)), len(read_graphdef(slim))))
x = torch.randn(4, 5)
out = TFNet.from_frozen_graph(path, ["x:0"], ["probs:0"])(x)
ref = torch.softmax(x @ torch.from_numpy(checkpoint["dense/kernel"]) + torch.from_numpy(checkpoint["dense/bias"]), -1)
print("frozen graph matches the reference computation:", bool(torch.allclose(out, ref, atol=1e-6)))
