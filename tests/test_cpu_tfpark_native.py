"""TFPark training / prediction classes with torch content (reference surface: pyzoo/zoo/tfpark/tf_optimizer.py:342-709,
estimator.py:30-332, tf_predictor.py:27-90), the GraphDef utilities (pyzoo/zoo/util/tf_graph_util.py) and the automl remote
helpers (pyzoo/zoo/automl/common/util.py:193-239)."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from zoo.common.triggers import MaxEpoch
from zoo.pipeline.api.keras.optimizers import Adam
from zoo.tfpark import TFDataset, TFEstimator, TFEstimatorSpec, TFOptimizer, TFPredictor, ZooOptimizer


def _data(n=512, seed=0):
    rng = np.random.RandomState(seed)
    x = rng.randn(n, 8).astype(np.float32)
    y = (x @ rng.randn(8, 3).astype(np.float32)).argmax(1).astype(np.int64)
    return x, y


def test_tfoptimizer_from_loss_trains_and_predicts(tmp_path):
    x, y = _data()
    net = nn.Sequential(nn.Linear(8, 32), nn.ReLU(), nn.Linear(32, 3))
    ds = TFDataset.from_ndarrays((x, y), batch_size=64, val_tensors=(x[:128], y[:128]))
    opt = TFOptimizer.from_loss(nn.CrossEntropyLoss(), ZooOptimizer(Adam(lr=1e-2)), model=net, dataset=ds, val_method=["accuracy"],
                                clip_norm=5.0, model_dir=str(tmp_path))
    opt.optimize(end_trigger=MaxEpoch(8))
    acc = (net.cpu()(torch.from_numpy(x)).argmax(1).numpy() == y).mean()
    assert acc > 0.9
    pred = TFPredictor.from_keras(net, TFDataset.from_ndarrays(x, batch_per_thread=100)).predict()
    assert len(pred) == len(x) and pred[0].shape == (3,)
    assert (np.stack(pred).argmax(1) == y).mean() == pytest.approx(acc, abs=1e-6)
    with pytest.raises(ValueError):
        TFPredictor.from_keras(net, TFDataset.from_ndarrays(x, batch_size=64))
    with pytest.raises(NotImplementedError):
        TFOptimizer.from_train_op(None, None)


def test_tfoptimizer_from_keras():
    from zoo.pipeline.api.keras.layers import Dense
    from zoo.pipeline.api.keras.models import Sequential

    x, y = _data(256, 1)
    m = Sequential()
    m.add(Dense(16, activation="relu", input_shape=(8,)))
    m.add(Dense(3))
    m.compile(optimizer=Adam(lr=1e-2), loss=nn.CrossEntropyLoss(), metrics=["accuracy"])
    TFOptimizer.from_keras(m, TFDataset.from_ndarrays((x, y), batch_size=64)).optimize(MaxEpoch(10))
    assert m.evaluate(x, y, batch_size=128)[1] > 0.8


def test_tfestimator_model_fn_state_lives_in_model_dir(tmp_path):
    x, y = _data()

    def model_fn(features, labels, mode, params):
        net = nn.Sequential(nn.Linear(features.shape[1], 32), nn.ReLU(), nn.Linear(32, params["classes"]))
        return TFEstimatorSpec(mode, model=net, loss=nn.CrossEntropyLoss(), train_op=ZooOptimizer(Adam(lr=1e-2)),
                               predictions=lambda o: {"cls": o.argmax(-1), "logits": o})

    def input_fn(mode):
        if mode == "train":
            return TFDataset.from_ndarrays((x, y), batch_size=64)
        if mode == "eval":
            return TFDataset.from_ndarrays((x, y), batch_per_thread=128)
        return TFDataset.from_ndarrays(x, batch_per_thread=128)

    est = TFEstimator.from_model_fn(model_fn, model_dir=str(tmp_path), params={"classes": 3})
    est.train(input_fn, steps=60)
    assert est.global_step == 60
    ev = est.evaluate(input_fn, ["acc", "loss"])
    assert ev["acc"] > 0.9 and ev["loss"] < 0.5
    other = TFEstimator.from_model_fn(model_fn, model_dir=str(tmp_path), params={"classes": 3})     # a fresh object sees the trained state
    pr = other.predict(input_fn, predict_keys=["cls"])
    assert len(pr) == len(x) and set(pr[0]) == {"cls"}
    assert (np.array([q["cls"] for q in pr]) == y).mean() == pytest.approx(ev["acc"], abs=1e-6)
    other.train(input_fn, steps=5)
    assert other.global_step == 65


def test_graph_utils_without_tensorflow(tmp_path):
    from zoo.pipeline.api.net import TFNet
    from zoo.pipeline.api.net.utils import find_placeholders, find_tensors
    from zoo.tfpark.graphdef import read_graphdef, write_graphdef
    from zoo.util import tf_graph_util as g
    from zoo.util.tf import export_tf

    net = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 2))
    export_tf(net, str(tmp_path), torch.randn(1, 4))
    pb = os.path.join(str(tmp_path), "frozen_inference_graph.pb")
    nodes = read_graphdef(pb)
    relu = next(n.name for n in nodes if n.op == "Relu")
    sub = read_graphdef(g.extract_sub_graph(pb, [relu]))
    assert sub[-1].name == relu and len(sub) < len(nodes)
    with pytest.raises(TypeError):
        g.extract_sub_graph(pb, relu)
    assert [n.name for n in find_placeholders(["output:0"], pb)] == ["input"]
    assert [n.op for n in find_tensors(["output"], lambda n: n.op == "MatMul", pb)] == ["MatMul", "MatMul"]
    w = next(n for n in nodes if n.op == "Const")
    assert g.tensor_shape_from_node_def_name(pb, w.name + ":0") == [4, 8] and not g.must_run_on_cpu(w)
    names = [n.name for n in read_graphdef(g.remove_training_nodes(pb))]
    assert "output" not in names
    assert "output" in [n.name for n in read_graphdef(g.remove_training_nodes(pb, protected_nodes=["output"]))]
    graph = write_graphdef(None, [("x", "Placeholder", [], {"dtype": np.float32}), ("w", "VariableV2", [], {"dtype": np.float32}),
                                  ("w/read", "Identity", ["w"], {"T": np.float32}),
                                  ("chk", "CheckNumerics", ["w/read"], {"T": np.float32}),
                                  ("y", "MatMul", ["x", "chk"], {"T": np.float32, "transpose_a": False, "transpose_b": False})])
    frozen = g.convert_variables_to_constants({"w": 2 * np.eye(3, dtype=np.float32)}, graph, ["y"])
    assert [(n.name, n.op) for n in read_graphdef(frozen)][1] == ("w", "Const")
    slim = g.remove_training_nodes(frozen)
    assert [n.op for n in read_graphdef(slim)] == ["Placeholder", "Const", "MatMul"]
    out = TFNet.from_frozen_graph(slim, ["x:0"], ["y:0"])(torch.ones(2, 3))
    assert torch.equal(out, 2 * torch.ones(2, 3))
    with pytest.raises(KeyError):
        g.convert_variables_to_constants({}, graph, ["y"])


def test_automl_remote_helpers(tmp_path):
    from zoo.automl.common.util import get_remote_list, restore_hdfs, save_zip, upload_ppl_hdfs

    class M:
        def save(self, mp, cp):
            open(mp, "w").write("weights")

        def restore(self, mp, **kw):
            self.got = open(mp).read()

    z = os.path.join(str(tmp_path), "best", "ppl.zip")
    save_zip(z, None, M(), {"a": 1})
    remote = "file://" + os.path.join(str(tmp_path), "remote")
    upload_ppl_hdfs(remote, z)
    assert get_remote_list(remote) == ["ppl.zip"]
    m = M()
    assert restore_hdfs(z, remote, None, m) == {"a": 1} and m.got == "weights"


def test_zoo_optimizer_wrapper_and_fake_optim_method():
    """ZooOptimizer hands the wrapped description to the engine unchanged; FakeOptimMethod leaves the weights where they are
    (reference: pyzoo/zoo/tfpark/zoo_optimizer.py:22-140)."""
    import numpy as np
    import torch

    from zoo.common.triggers import MaxIteration
    from zoo.feature.common import FeatureSet
    from zoo.pipeline.api.keras.optimizers import Adam, to_opt_config
    from zoo.pipeline.estimator import Estimator
    from zoo.tfpark.unsupported import ZooOptimizer as FromOldHome
    from zoo.tfpark.zoo_optimizer import FakeOptimMethod, ZooOptimizer

    assert FromOldHome is ZooOptimizer
    wrapped = ZooOptimizer(ZooOptimizer(Adam(lr=0.02)))
    cfg, ref = to_opt_config(wrapped), to_opt_config(Adam(lr=0.02))
    assert (cfg.name, cfg.lr) == (ref.name, ref.lr) and wrapped.lr == 0.02 and wrapped.name == "Adam"
    x, y = np.random.RandomState(0).randn(16, 4).astype(np.float32), np.zeros((16, 1), np.float32)
    m = torch.nn.Linear(4, 1)
    before = [p.detach().clone() for p in m.parameters()]
    Estimator(m, FakeOptimMethod()).engine.train(FeatureSet.numpy(x, y), "mse", MaxIteration(2), batch_size=8)
    assert all(torch.equal(a, b) for a, b in zip(m.parameters(), before))


def test_dataset_like_iterables_stand_in_for_tf_data():
    """from_tf_data_dataset / FeatureSet.tf_dataset take what a tf.data pipeline is to its consumer: an iterable of elements."""
    import numpy as np

    from zoo.feature.common import FeatureSet
    from zoo.tfpark import TFDataset

    rng = np.random.RandomState(0)
    elems = [((rng.rand(4).astype(np.float32), rng.rand(2).astype(np.float32)), np.int64(i % 3)) for i in range(10)]
    ds = TFDataset.from_tf_data_dataset(elems, batch_size=5, validation_dataset=elems[:4])
    b = next(iter(ds.get_training_data().batches(5, train=False)))
    assert [tuple(t.shape) for t in b.inputs] == [(5, 4), (5, 2)] and b.targets.flatten().tolist() == [0, 1, 2, 0, 1]
    assert ds.get_validation_data().size() == 4

    class Pipeline:                                    # quacks like tf.data.Dataset
        def as_numpy_iterator(self):
            return iter(np.arange(12, dtype=np.float32).reshape(6, 2))

    fs = FeatureSet.tf_dataset(lambda: Pipeline(), total_size=4)
    assert fs.size() == 4
    only_x = TFDataset.from_tf_data_dataset(Pipeline(), batch_per_thread=3)
    assert only_x.get_training_data().size() == 6
