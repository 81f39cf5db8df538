"""NNFrames on pandas DataFrames, TorchNet / TorchCriterion gradient match, Net loaders, GraphNet, JTensor
(reference tests: pyzoo/test/zoo/pipeline/nnframes/test_nn_classifier.py, pyzoo/test/zoo/pipeline/api/test_torch_net.py)."""
import numpy as np
import pandas as pd
import pytest
import torch
import torch.nn as nn


@pytest.fixture(autouse=True)
def _cpu_ctx(monkeypatch):
    import zoo.common.nncontext as nc

    monkeypatch.setenv("ZOO_FORCE_CPU", "1")
    nc._CONTEXT = None
    yield
    nc._CONTEXT = None


def _df(n=400, seed=0):
    rng = np.random.RandomState(seed)
    x = rng.randn(n, 4).astype(np.float32)
    y = (x.sum(1) > 0).astype(np.float32) + 1.0          # 1-based labels
    return pd.DataFrame({"features": list(x), "label": y, "id": np.arange(n)})


def test_nnestimator_params_and_fit_transform():
    from zoo.common.triggers import EveryEpoch
    from zoo.feature.common import SeqToTensor
    from zoo.pipeline.api.keras.objectives import MeanSquaredError
    from zoo.pipeline.api.keras.optimizers import Adam
    from zoo.pipeline.nnframes import NNEstimator, NNModel

    torch.manual_seed(0)
    model = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 1))
    est = NNEstimator(model, MeanSquaredError(), SeqToTensor([4]), SeqToTensor([1]))
    assert est.getBatchSize() == 1 and est.getMaxEpoch() == 50 and est.getLearningRate() == 1e-3 and est.isCachingSample()
    est.setBatchSize(40).setMaxEpoch(8).setOptimMethod(Adam(lr=1e-2)).setLabelCol("label").setFeaturesCol("features").setPredictionCol("pred")
    est.setConstantGradientClipping(-5, 5).setLearningRateDecay(0.0).setCachingSample(False)
    df = _df()
    est.setValidation(EveryEpoch(), df, ["mae"], 40)
    m = est.fit(df)
    assert isinstance(m, NNModel) and m.getPredictionCol() == "pred" and m.getBatchSize() == 40
    out = m.transform(df)
    assert list(out.columns) == ["features", "label", "id", "pred"] and len(out) == len(df)
    pred = np.asarray([p[0] for p in out["pred"]])
    assert np.mean((pred - df["label"].values) ** 2) < 0.2
    with pytest.raises(ValueError):
        est.setDataCacheLevel("DISK_AND_DRAM")
    est.setDataCacheLevel("DISK_AND_DRAM", 2)
    assert est.getDataCacheLevel() == ("DISK_AND_DRAM", 2)


def test_nnclassifier_and_persistence(tmp_path):
    from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
    from zoo.pipeline.api.keras.optimizers import Adam
    from zoo.pipeline.nnframes import NNClassifier, NNClassifierModel, NNModel

    torch.manual_seed(0)
    model = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 2), nn.Softmax(dim=-1))
    clf = NNClassifier(model, SparseCategoricalCrossEntropy(zero_based_label=False), [4]).setBatchSize(50).setMaxEpoch(12).setOptimMethod(Adam(lr=2e-2))
    df = _df()
    m = clf.fit(df)
    assert isinstance(m, NNClassifierModel)
    out = m.transform(df)
    assert out["prediction"].dtype == np.float64 and set(out["prediction"]) <= {1.0, 2.0}
    assert (out["prediction"] == df["label"]).mean() > 0.9
    m.write().overwrite().save(str(tmp_path / "nnmodel"))
    import json
    import os

    meta = json.loads(open(tmp_path / "nnmodel" / "metadata" / "part-00000").readline())
    assert meta["tensorDataType"] == "TensorFloat" and meta["modelType"] == "default" and meta["paramMap"]["batchSize"] == 50
    assert {"module", "weight", "samplePreprocessing", "metadata"} <= set(os.listdir(tmp_path / "nnmodel"))
    m2 = NNModel.load(str(tmp_path / "nnmodel"))
    assert isinstance(m2, NNClassifierModel)
    assert (m2.transform(df)["prediction"] == out["prediction"]).all()
    binary = NNClassifierModel(nn.Sequential(nn.Linear(4, 1), nn.Sigmoid()), [4]).setThreshold(0.5)
    assert set(binary.transform(df)["prediction"]) <= {0.0, 1.0}


def test_torchnet_torchcriterion_gradient_match(tmp_path):
    """Same input through plain torch and through TorchNet / TorchCriterion inside the engine (test_torch_net.py:124-153)."""
    from zoo.feature.common import MiniBatch
    from zoo.pipeline.api.net import Net, TorchCriterion, TorchNet
    from zoo.pipeline.engine import TrainingEngine

    class LeNetish(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv = nn.Conv2d(1, 4, 3)
            self.fc = nn.Linear(4 * 6 * 6, 3)

        def forward(self, x):
            return torch.log_softmax(self.fc(torch.relu(self.conv(x)).flatten(1)), -1)

    torch.manual_seed(0)
    ref = LeNetish()
    import copy

    net = TorchNet.from_pytorch(copy.deepcopy(ref), [2, 1, 8, 8])
    assert net.traced is not None
    crit = TorchCriterion.from_pytorch(nn.NLLLoss(), [1, 3], torch.LongTensor([1]))
    x, y = torch.randn(6, 1, 8, 8), torch.randint(0, 3, (6,))
    loss_ref = nn.NLLLoss()(ref(x), y)
    loss_ref.backward()
    eng = TrainingEngine(net, torch.optim.SGD(net.parameters(), lr=0.0)).prepare()
    loss, _ = eng.forward_loss(crit, x, y)
    loss.backward()
    assert abs(float(loss) - float(loss_ref)) < 1e-6
    arena = eng.fopt.arena                      # gradients land in the flat arena (every tensor starts on an aligned offset)
    for p, q in zip(net.module.parameters(), ref.parameters()):
        a, b = arena.range_of(p)
        torch.testing.assert_close(arena.flat_g[a:b].view(q.shape), q.grad, atol=1e-6, rtol=1e-5)
    np.testing.assert_allclose(net.get_weights(), torch.cat([p.detach().reshape(-1) for p in ref.parameters()]).numpy())
    net.savePytorch(str(tmp_path / "m.pt"))
    loaded = Net.load_torch(str(tmp_path / "m.pt"))
    torch.testing.assert_close(loaded(x), ref(x))
    assert loaded.predict(x.numpy(), batch_size=4).shape == (6, 3)
    w = net.get_weights()
    net.set_weights(np.zeros_like(w))
    assert float(net(x).exp().sum()) == pytest.approx(6.0)
    # Keras-1 JSON definition + weights (npz) -> zoo Sequential
    import json

    spec = {"class_name": "Sequential", "config": [
        {"class_name": "Dense", "config": {"output_dim": 5, "activation": "relu", "batch_input_shape": [None, 4], "W_regularizer": None}},
        {"class_name": "Dropout", "config": {"p": 0.2}},
        {"class_name": "Dense", "config": {"output_dim": 3, "activation": "softmax"}}]}
    (tmp_path / "k.json").write_text(json.dumps(spec))
    rng = np.random.RandomState(0)
    ws = [rng.randn(4, 5).astype(np.float32), rng.randn(5).astype(np.float32), rng.randn(5, 3).astype(np.float32), rng.randn(3).astype(np.float32)]
    np.savez(str(tmp_path / "k.npz"), *ws)
    km = Net.load_keras(str(tmp_path / "k.json"), str(tmp_path / "k.npz")).eval()
    xk = torch.randn(7, 4)
    want = torch.softmax(torch.relu(xk @ torch.from_numpy(ws[0]) + torch.from_numpy(ws[1])) @ torch.from_numpy(ws[2]) + torch.from_numpy(ws[3]), -1)
    torch.testing.assert_close(km(xk), want, atol=1e-5, rtol=1e-5)
    with pytest.raises(NotImplementedError):
        Net.load_keras(None, "weights.h5")


def test_graphnet_and_multi_io():
    from zoo.pipeline.api.net import GraphNet, TorchNet

    torch.manual_seed(0)
    m = nn.Sequential()
    m.add_module("fc1", nn.Linear(4, 6)); m.add_module("act", nn.ReLU()); m.add_module("fc2", nn.Linear(6, 2))
    g = GraphNet(m)
    x = torch.randn(3, 4)
    torch.testing.assert_close(g.new_graph("act")(x), torch.relu(m.fc1(x)))
    a, b = g.new_graph(["fc1", "fc2"])(x)
    assert a.shape == (3, 6) and b.shape == (3, 2)
    g.freeze_up_to(["fc1"])
    assert not m.fc1.weight.requires_grad and m.fc2.weight.requires_grad
    g.unfreeze()
    assert m.fc1.weight.requires_grad and len(g.layers) == 3
    with pytest.raises(KeyError):
        g.new_graph("nope")

    class TwoInTwoOut(nn.Module):
        def forward(self, a, b):
            return a + b, a * b

    t = TorchNet.from_pytorch(TwoInTwoOut(), [[2, 3], [2, 3]])
    o1, o2 = t(torch.ones(2, 3), torch.full((2, 3), 2.0))
    assert float(o1.sum()) == 18 and float(o2.sum()) == 12


def test_jtensor_and_callzoofunc():
    from zoo.common.utils import JTensor, callZooFunc, register_zoo_func

    a = np.arange(6, dtype=np.float32).reshape(2, 3)
    jt = JTensor.from_ndarray(a)
    np.testing.assert_array_equal(jt.to_ndarray(), a)
    sp = JTensor.sparse(np.array([1.0, 2.0]), np.array([[0, 1], [2, 0]]), (2, 3))
    np.testing.assert_array_equal(sp.to_ndarray(), np.array([[0, 0, 1.0], [2.0, 0, 0]]))
    assert sp.to_tensor().is_sparse and JTensor.from_ndarray(None) is None
    register_zoo_func("myAdd", lambda a, b: a + b)
    assert callZooFunc("float", "myAdd", 2, 3) == 5
    with pytest.raises(Exception):
        callZooFunc("float", "nope")


def test_nnimage_reader(tmp_path):
    import cv2

    from zoo.pipeline.nnframes import NNImageReader

    for i in range(3):
        cv2.imwrite(str(tmp_path / f"im{i}.png"), np.full((10 + i, 12, 3), i * 40, np.uint8))
    df = NNImageReader.readImages(str(tmp_path), resizeH=8, resizeW=6)
    assert len(df) == 3
    r = df["image"][0]
    assert (r["height"], r["width"], r["nChannels"]) == (8, 6, 3) and r["origin"].startswith("file:") and r["data"].shape == (8, 6, 3)


def test_pojo_inference_api(tmp_path):
    """AbstractInferenceModel / JTensor / InferenceModelFactory / FloatModel (reference: pipeline/inference/*.java,
    InferenceModelFactory.scala, FloatModel.scala): the three predict forms, reload, weight-sharing copies."""
    import torch
    import torch.nn as nn

    from zoo.pipeline.inference import AbstractInferenceModel, FloatModel, InferenceModelFactory, InferenceRuntimeException, JTensor

    torch.manual_seed(0)
    net = nn.Sequential(nn.Linear(4, 3), nn.ReLU(), nn.Linear(3, 2)).eval()
    path = str(tmp_path / "model.pt")
    torch.save(net, path)

    class TextModel(AbstractInferenceModel):                 # "extend AbstractInferenceModel", as the programming guide says
        pass

    m = TextModel(2)
    with pytest.raises(InferenceRuntimeException):
        m.predict([1.0, 2.0, 3.0, 4.0], 1, 4)
    m.loadPyTorch(path)
    x = np.arange(8, dtype=np.float32).reshape(2, 4)
    want = net(torch.from_numpy(x)).detach().numpy()
    flat = m.predict(x.reshape(-1).tolist(), 2, 4)
    assert isinstance(flat, list) and np.allclose(flat, want.reshape(-1), atol=1e-5)
    recs = m.predict([[JTensor.from_ndarray(x[0])], [JTensor.from_ndarray(x[1])]])
    assert len(recs) == 2 and len(recs[0]) == 1 and list(recs[0][0].shape) == [2] and np.allclose(recs[1][0].to_ndarray(), want[1], atol=1e-5)
    assert np.allclose(m.predict(x), want, atol=1e-5)
    with open(path, "rb") as f:
        m2 = AbstractInferenceModel(True, 3)                  # (autoScalingEnabled, concurrentNum)
        m2.loadPyTorch(f.read())
    assert m2.auto_scaling_enabled and m2.concurrent_num == 3 and np.allclose(m2.predict(x), want, atol=1e-5)
    torch.save(nn.Sequential(nn.Linear(4, 2)).eval(), str(tmp_path / "other.pt"))
    m.reload(str(tmp_path / "other.pt"))
    assert np.asarray(m.predict(x)).shape == (2, 2) and not np.allclose(m.predict(x), want, atol=1e-3)

    fm = InferenceModelFactory.load_float_model_for_pytorch(path)
    assert isinstance(fm, FloatModel) and np.allclose(fm.predict(x).numpy(), want, atol=1e-5)
    clones = fm.copy(2)
    assert len(clones) == 2 and clones[0].model is not fm.model and not clones[0].is_original
    assert next(clones[0].model.parameters()).data_ptr() == next(fm.model.parameters()).data_ptr()          # weights shared
    out = clones[1].predict([[JTensor.from_ndarray(x[0])], [JTensor.from_ndarray(x[1])]])
    assert np.allclose(out[0][0].to_ndarray(), want[0], atol=1e-5)
    fm.release()
    assert fm.isReleased()
