"""BigDL model files (`Module.saveModule` / `ZooModel.saveModel`) <-> torch: writer / loader round trip, `Net.load_bigdl`,
`InferenceModel.load`, and -- when the reference tree is on the machine -- the reference's own saved models
(zoo/src/test/resources/models/bigdl/bigdl_lenet.model, models/zoo_keras/*.model; NetSpec.scala:113-123 loads the same files)."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

REF = "/root/reference/zoo/src/test/resources/models"


def _net():
    torch.manual_seed(0)
    net = nn.Sequential(nn.Conv2d(3, 8, 3, stride=2, padding=1), nn.BatchNorm2d(8), nn.ReLU(), nn.MaxPool2d(2), nn.Conv2d(8, 8, 3, groups=2), nn.Tanh(),
                        nn.AvgPool2d(2), nn.Flatten(), nn.Linear(8, 16), nn.ELU(0.7), nn.Dropout(0.3), nn.Sequential(nn.Linear(16, 4),
                                nn.LogSoftmax(-1))).eval()
    with torch.no_grad():
        net[1].running_mean.uniform_(-1, 1)
        net[1].running_var.uniform_(0.5, 2)
    return net


def test_bigdl_round_trip_and_public_loaders(tmp_path):
    from zoo.models.bigdl import BigDLConversionException, load_bigdl, save_bigdl
    from zoo.pipeline.api.net import Net
    from zoo.pipeline.inference import InferenceModel
    from zoo.util import protowire as pw

    net = _net()
    path = str(tmp_path / "model.bigdl")
    buf = save_bigdl(net, path)
    assert open(path, "rb").read() == buf
    # the file is BigDL's layout: module type names of the JVM classes, parameters in the root's global storage
    assert b"com.intel.analytics.bigdl.nn.SpatialConvolution" in buf and b"global_storage" in buf
    x = torch.randn(5, 3, 20, 20)
    want = net(x)
    back = load_bigdl(path)
    assert not back.training
    torch.testing.assert_close(back(x), want, atol=0, rtol=0)
    assert [type(m).__name__ for m in back][:4] == ["Conv2d", "BatchNorm2d", "ReLU", "MaxPool2d"]
    g = Net.load_bigdl(path)
    torch.testing.assert_close(g.forward(x), want)
    torch.testing.assert_close(Net.load(path).forward(x), want)
    im = InferenceModel(1, "cpu")
    im.load(path)
    np.testing.assert_allclose(im.predict(x.numpy()), want.detach().numpy(), atol=1e-6)
    assert Net.save_bigdl(net, str(tmp_path / "again.bigdl")) and os.path.getsize(str(tmp_path / "again.bigdl")) == len(buf)
    with pytest.raises(NotImplementedError):
        save_bigdl(nn.Sequential(nn.LSTM(4, 4)))
    with pytest.raises(BigDLConversionException):
        load_bigdl(pw.encode({"moduleType": "com.intel.analytics.bigdl.nn.NoSuchLayer", "subModules": []},
                             __import__("zoo.models.bigdl.bigdl_schema", fromlist=["MODULE"]).MODULE))
    with pytest.raises(BigDLConversionException):
        load_bigdl(b"\x0a\x03abc")


def test_bigdl_graph_and_table_layers():
    """A StaticGraph with a residual branch (CAddTable) written by hand in the wire format."""
    from zoo.models.bigdl import bigdl_schema as S
    from zoo.models.bigdl import load_bigdl
    from zoo.models.bigdl.bigdl_writer import _Writer
    from zoo.util import protowire as pw

    torch.manual_seed(1)
    lin = nn.Linear(6, 6)
    w = _Writer()
    nodes = [w.module("Input", "in"), dict(w.convert(lin, "fc"), preModules=["in"]), dict(w.module("ReLU", "act"), preModules=["fc"]),
             dict(w.module("CAddTable", "sum"), preModules=["in", "act"]), dict(w.module("SoftMax", "prob"), preModules=["sum"])]
    root = w.module("StaticGraph", "g", subs=nodes)
    root["attr"] = [{"key": "inputNames", "value": {"dataType": S.DT_ARRAY, "arrayValue": {"size": 1, "datatype": S.DT_STRING, "str": ["in"]}}},
                    {"key": "outputNames", "value": {"dataType": S.DT_ARRAY, "arrayValue": {"size": 1, "datatype": S.DT_STRING, "str": ["prob"]}}},
                    {"key": "global_storage", "value": {"nameAttrListValue": {"name": "global_storage", "attr": w.storage_entries}}}]
    net = load_bigdl(pw.encode(root, S.MODULE))
    x = torch.randn(3, 6)
    torch.testing.assert_close(net(x), torch.softmax(x + torch.relu(lin(x)), -1))
    assert net.order == ["in", "fc", "act", "sum", "prob"]


@pytest.mark.skipif(not os.path.exists(REF + "/bigdl/bigdl_lenet.model"), reason="reference test resources are not on this machine")
def test_reference_saved_models_load_and_run():
    from zoo.models.bigdl import load_bigdl

    lenet = load_bigdl(REF + "/bigdl/bigdl_lenet.model")
    assert lenet.order[0] == "reshape1" and lenet.order[-1] == "logSoftMax"
    y = lenet(torch.rand(4, 1, 28, 28))
    assert y.shape == (4, 5)
    torch.testing.assert_close(y.exp().sum(1), torch.ones(4))                       # a log-softmax over 5 classes
    conv = lenet.nodes["conv1_5x5"]
    assert conv.weight.shape == (6, 1, 5, 5) and float(conv.weight.abs().sum()) > 0  # the stored parameters arrived
    # the two Keras-style models of NetSpec "net load model": forward on Tensor(2, 3, 5)
    assert load_bigdl(REF + "/zoo_keras/small_seq.model")(torch.rand(2, 3, 5)).numel() == 30
    assert load_bigdl(REF + "/zoo_keras/small_model.model")(torch.rand(2, 3, 5)).shape == (2, 3, 7)


def test_separate_weight_file_and_reference_nnmodel_directory(tmp_path):
    """`saveModule(path, weightPath)`: definition and weights apart (weight file: big-endian records keyed by storage id + MD5),
    and the NNModel directory the reference's NNEstimatorSpec.scala:467-485 loads (models/nnframes_keras: metadata JSON + module +
    weight + a Java-serialized samplePreprocessing)."""
    import pandas as pd

    from zoo.models.bigdl import load_bigdl, save_bigdl
    from zoo.models.bigdl.bigdl_loader import read_weight_file
    from zoo.pipeline.api.net import Net
    from zoo.pipeline.nnframes import NNModel

    torch.manual_seed(0)
    net = nn.Sequential(nn.Linear(6, 5), nn.ReLU(), nn.Linear(5, 2)).eval()
    d, w = str(tmp_path / "m.bigdl"), str(tmp_path / "m.weight")
    save_bigdl(net, d, w)
    assert len(open(d, "rb").read()) < len(save_bigdl(net)) - 4 * 40              # the definition carries no tensor data
    ws = read_weight_file(w)
    assert sorted(v.size for v in ws.values()) == [2, 5, 10, 30]
    x = torch.randn(3, 6)
    assert torch.allclose(load_bigdl(d, w)(x), net(x), atol=1e-6)
    assert torch.allclose(Net.load_bigdl(d, w)(x), net(x), atol=1e-6)
    with pytest.raises(Exception):
        load_bigdl(d)                                                              # storages missing without the weight file

    ref = "/root/reference/zoo/src/test/resources/models/nnframes_keras"
    if not os.path.exists(ref):
        pytest.skip("reference test resources are not on this machine")
    ws = read_weight_file(ref + "/weight")
    assert sorted(v.size for v in ws.values()) == [2, 12]
    m = NNModel.load(ref)
    assert m.getBatchSize() == 10 and m.getFeaturesCol() == "features" and m.getPredictionCol() == "prediction"
    kernel = next(v for v in ws.values() if v.size == 12).reshape(2, 6)           # Dense(2) on Input(6): BigDL Linear weight [out, in]
    df = pd.DataFrame({"features": [list(map(float, range(6))), [1.0] * 6]})
    out = m.setBatchSize(8).transform(df)
    want = np.asarray(df["features"].tolist(), np.float32) @ kernel.T
    assert np.allclose(np.asarray(out["prediction"].tolist(), np.float32), want, atol=1e-5)
