"""TFPark facade (TFDataset constructors, KerasModel fit/evaluate/predict/save_weights), LocalEstimator and the
ONNX graph interpreter on hand-built GraphProto-shaped objects (reference tests: pyzoo/test/zoo/tfpark/*,
pyzoo/test/zoo/pipeline/onnx/*)."""
from types import SimpleNamespace as NS

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


def _xy(n=600):
    rng = np.random.RandomState(0)
    x = rng.randn(n, 6).astype(np.float32)
    return x, (x[:, :3].sum(1) > x[:, 3:].sum(1)).astype(np.int64)


def test_tfdataset_and_keras_model(tmp_path):
    from zoo.pipeline.api.keras.engine.topology import Sequential
    from zoo.pipeline.api.keras.layers import Dense
    from zoo.tfpark import KerasModel, TFDataset, TFEstimator, TFNet, TFOptimizer, ZooOptimizer

    x, y = _xy()
    torch.manual_seed(0)
    with pytest.raises(ValueError):
        TFDataset.from_ndarrays((x, y), batch_size=8, batch_per_thread=8)
    ds = TFDataset.from_ndarrays((x, y), batch_size=60, val_tensors=(x[:100], y[:100]))
    assert len(ds) == 600 and ds.get_validation_data().size() == 100
    rdd = TFDataset.from_rdd([(x[i], y[i]) for i in range(50)], batch_size=10)
    assert rdd.get_training_data().size() == 50
    df = pd.DataFrame({"f": list(x[:40]), "l": y[:40]})
    assert len(TFDataset.from_dataframe(df, ["f"], ["l"], batch_size=8)) == 40
    with pytest.raises(FileNotFoundError):      # the reader is real (zoo/tfpark/tfrecord.py); a missing file is an IO error
        TFDataset.from_tfrecord_file(str(tmp_path / "missing.tfrecord"))
    net = Sequential([Dense(16, activation="relu", input_shape=(6,)), Dense(2, activation="softmax")])
    net.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    km = KerasModel(net)
    km.fit(ds, epochs=25)
    loss, acc = km.evaluate(ds)
    assert acc > 0.9 and km.metrics_names == ["loss", "Top1Accuracy"]
    p = km.predict(TFDataset.from_ndarrays(x[:9], batch_per_thread=4))
    assert p.shape == (9, 2)
    km.save_weights(str(tmp_path / "w.pt"))
    net2 = Sequential([Dense(16, activation="relu", input_shape=(6,)), Dense(2, activation="softmax")])
    km2 = KerasModel(net2, optimizer="sgd", loss="sparse_categorical_crossentropy")
    km2.load_weights(str(tmp_path / "w.pt"))
    np.testing.assert_allclose(km2.predict(x[:9]), p, atol=1e-6)
    plain = KerasModel(nn.Sequential(nn.Linear(6, 2), nn.Softmax(-1)), optimizer="adam", loss="sparse_categorical_crossentropy")
    plain.fit(x, y, batch_size=100, epochs=1, distributed=True)
    # TFNet / TFOptimizer / TFEstimator are real (tests/test_cpu_tfnet.py, tests/test_cpu_tfpark_native.py); what needs a live
    # TensorFlow object says so
    with pytest.raises(NotImplementedError):
        TFOptimizer.from_loss(None, None)
    with pytest.raises(NotImplementedError):
        TFOptimizer.from_train_op(None, None)
    with pytest.raises(TypeError):
        TFEstimator(object())
    assert ZooOptimizer(torch.optim.SGD([torch.zeros(1, requires_grad=True)], lr=0.5)).defaults["lr"] == 0.5


def test_local_estimator():
    from zoo.pipeline.api.keras.optimizers import Adam
    from zoo.pipeline.estimator import LocalEstimator

    x, y = _xy(400)
    data = [(x[i], y[i]) for i in range(400)]
    est = LocalEstimator(nn.Sequential(nn.Linear(6, 16), nn.ReLU(), nn.Linear(16, 2), nn.Softmax(-1)), "sparse_categorical_crossentropy",
                         Adam(lr=1e-2), ["accuracy"], thread_num=4)
    est.fit(data, data[:100], batch_size=40, epoch=15)
    assert est.validate(data, batch_size=100)[0].value > 0.9


def _t(name, arr):
    return NS(name=name, numpy_value=np.asarray(arr))


def _node(op, ins, outs, **attrs):
    return NS(op_type=op, input=ins, output=outs, attribute=list(attrs.items()))


def test_onnx_graph_mlp_and_convnet():
    from zoo.pipeline.api.onnx import OnnxLoader

    rng = np.random.RandomState(0)
    w1, b1, w2, b2 = rng.randn(8, 5).astype(np.float32), rng.randn(8).astype(np.float32), rng.randn(8, 3).astype(np.float32), rng.randn(3).astype(np.float32)
    g = NS(initializer=[_t("w1", w1), _t("b1", b1), _t("w2", w2), _t("b2", b2)],
           input=[NS(name="x"), NS(name="w1"), NS(name="b1"), NS(name="w2"), NS(name="b2")],
           output=[NS(name="y")],
           node=[_node("Gemm", ["x", "w1", "b1"], ["h"], transB=1), _node("Relu", ["h"], ["r"]), _node("MatMul", ["r", "w2"], ["m"]),
                 _node("Add", ["m", "b2"], ["a"]), _node("Softmax", ["a"], ["y"], axis=1)])
    model = OnnxLoader(NS(graph=g)).to_keras()
    x = torch.randn(4, 5)
    ref = torch.softmax(torch.relu(x @ torch.from_numpy(w1).t() + torch.from_numpy(b1)) @ torch.from_numpy(w2) + torch.from_numpy(b2), 1)
    torch.testing.assert_close(model(x), ref)
    assert len(list(model.parameters())) == 4                            # initializers are trainable
    model(x).sum().backward()
    cw, cb = rng.randn(4, 3, 3, 3).astype(np.float32), rng.randn(4).astype(np.float32)
    sc, bi, mu, va = np.ones(4, np.float32), np.zeros(4, np.float32), rng.randn(4).astype(np.float32), np.abs(rng.randn(4)).astype(np.float32) + 0.5
    g2 = NS(initializer=[_t("cw", cw), _t("cb", cb), _t("sc", sc), _t("bi", bi), _t("mu", mu), _t("va", va), _t("shape", np.array([0, -1], np.int64))],
            input=[NS(name="img")], output=[NS(name="out")],
            node=[_node("Conv", ["img", "cw", "cb"], ["c"], kernel_shape=[3, 3], pads=[1, 1, 1, 1], strides=[1, 1]),
                  _node("BatchNormalization", ["c", "sc", "bi", "mu", "va"], ["bn"], epsilon=1e-5), _node("LeakyRelu", ["bn"], ["l"], alpha=0.1),
                  _node("MaxPool", ["l"], ["p"], kernel_shape=[2, 2], strides=[2, 2]), _node("GlobalAveragePool", ["p"], ["gap"]),
                  _node("Reshape", ["gap", "shape"], ["out"])])
    m2 = OnnxLoader(g2).to_keras()
    img = torch.randn(2, 3, 8, 8)
    c = torch.nn.functional.conv2d(img, torch.from_numpy(cw), torch.from_numpy(cb), padding=1)
    bn = (c - torch.from_numpy(mu).view(1, -1, 1, 1)) / torch.sqrt(torch.from_numpy(va).view(1, -1, 1, 1) + 1e-5)
    want = torch.nn.functional.max_pool2d(torch.nn.functional.leaky_relu(bn, 0.1), 2).mean((2, 3))
    torch.testing.assert_close(m2(img), want, atol=1e-5, rtol=1e-4)
    with pytest.raises(NotImplementedError):
        OnnxLoader(NS(initializer=[], input=[NS(name="x")], output=[NS(name="y")], node=[_node("NonMaxSuppression", ["x"], ["y"])])).to_keras()
    with pytest.raises(ImportError):
        OnnxLoader.from_path("model.onnx")


@pytest.mark.parametrize("op,ins,attrs,ref", [
    ("Abs", [np.array([-1.0, 2.0])], {}, lambda a: np.abs(a)),
    ("Clip", [np.array([-2.0, 0.5, 3.0])], {"min": -1.0, "max": 1.0}, lambda a: np.clip(a, -1, 1)),
    ("Concat", [np.ones((2, 2)), np.zeros((2, 3))], {"axis": 1}, lambda a, b: np.concatenate([a, b], 1)),
    ("Transpose", [np.arange(6.0).reshape(2, 3)], {"perm": [1, 0]}, lambda a: a.T),
    ("Unsqueeze", [np.arange(3.0)], {"axes": [0]}, lambda a: a[None]),
    ("Squeeze", [np.zeros((1, 3, 1))], {"axes": [0, 2]}, lambda a: a.reshape(3)),
    ("Gather", [np.arange(12.0).reshape(4, 3), np.array([3, 0])], {"axis": 0}, lambda a, i: a[i]),
    ("Slice", [np.arange(20.0).reshape(4, 5)], {"starts": [1, 0], "ends": [3, 4], "axes": [0, 1]}, lambda a: a[1:3, 0:4]),
    ("ReduceMean", [np.arange(6.0).reshape(2, 3)], {"axes": [1], "keepdims": 0}, lambda a: a.mean(1)),
    ("ReduceSum", [np.arange(6.0).reshape(2, 3)], {"axes": [0], "keepdims": 1}, lambda a: a.sum(0, keepdims=True)),
    ("HardSigmoid", [np.array([-5.0, 0.0, 5.0])], {"alpha": 0.2, "beta": 0.5}, lambda a: np.clip(0.2 * a + 0.5, 0, 1)),
    ("Flatten", [np.zeros((2, 3, 4))], {"axis": 1}, lambda a: a.reshape(2, 12)),
    ("Pow", [np.array([2.0, 3.0]), np.array([2.0, 2.0])], {}, lambda a, b: a ** b),
    ("AveragePool", [np.arange(16.0).reshape(1, 1, 4, 4)], {"kernel_shape": [2, 2], "strides": [2, 2]}, lambda a: a.reshape(1, 1, 2, 2, 2, 2).mean((3, 5))),
    ("Shape", [np.zeros((2, 5))], {}, lambda a: np.array([2, 5])),
    ("LogSoftmax", [np.array([[1.0, 2.0, 3.0]])], {"axis": 1}, lambda a: a - np.log(np.exp(a).sum(1, keepdims=True))),
])
def test_onnx_run_node(op, ins, attrs, ref):
    from zoo.pipeline.api.onnx import OnnxLoader

    out = OnnxLoader.run_node(_node(op, [f"i{k}" for k in range(len(ins))], ["o"], **attrs), ins)[0]
    np.testing.assert_allclose(out, ref(*ins), atol=1e-6)


def test_tfrecord_examples_roundtrip_and_dataset(tmp_path):
    import struct

    from zoo.tfpark import TFDataset
    from zoo.tfpark.tfrecord import decode_example, encode_example, read_examples, read_tfrecord, write_tfrecord

    rng = np.random.RandomState(0)
    recs = [{"x": rng.randn(6).astype(np.float32), "label": np.int64(i % 3), "ids": np.array([-1, 2 ** 40, i], np.int64), "name": f"r{i}",
             "blob": [b"\\x00\\x01", b""]} for i in range(25)]
    ex = decode_example(encode_example(recs[3]))
    np.testing.assert_allclose(ex["x"], recs[3]["x"])
    assert ex["label"].tolist() == [0] and ex["ids"].tolist() == [-1, 2 ** 40, 3] and ex["name"] == [b"r3"] and ex["blob"] == [b"\\x00\\x01", b""]
    path = str(tmp_path / "data.tfrecord")
    assert write_tfrecord(path, recs) == 25
    assert write_tfrecord(path, [b"raw-payload"], append=True) == 1
    payloads = list(read_tfrecord(path))
    assert len(payloads) == 26 and payloads[-1] == b"raw-payload"
    # the framing is TensorFlow's: u64 length, masked crc of the length, payload, masked crc of the payload
    raw = open(path, "rb").read()
    assert struct.unpack("<Q", raw[:8])[0] == len(payloads[0])
    # corruption is detected
    bad = bytearray(raw); bad[20] ^= 0xFF
    (tmp_path / "bad.tfrecord").write_bytes(bytes(bad))
    with pytest.raises(IOError):
        list(read_tfrecord(str(tmp_path / "bad.tfrecord")))
    (tmp_path / "short.tfrecord").write_bytes(raw[:-3])
    with pytest.raises(IOError):
        list(read_tfrecord(str(tmp_path / "short.tfrecord")))
    write_tfrecord(str(tmp_path / "train.tfrecord"), recs)
    ds = TFDataset.from_tfrecord_file(None, str(tmp_path / "train.tfrecord"), batch_size=5, features="x", labels="label")
    assert len(ds) == 25
    xs = np.stack([e["x"] for e in read_examples(str(tmp_path / "train*.tfrecord"))])
    np.testing.assert_allclose(xs, np.stack([r["x"] for r in recs]))
    m = nn.Sequential(nn.Linear(6, 3))
    from zoo.tfpark import KerasModel

    km = KerasModel(m, optimizer="adam", loss="sparse_categorical_crossentropy")
    km.fit(ds, epochs=1)


def test_example_wire_format_matches_protobuf():
    """The hand-written Example codec against google.protobuf with the tf.train.Example schema built at run time."""
    pb = pytest.importorskip("google.protobuf")
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    from zoo.tfpark.tfrecord import decode_example, encode_example

    f = descriptor_pb2.FileDescriptorProto(name="zoo_example_test.proto", package="zoo_t", syntax="proto3")
    T = descriptor_pb2.FieldDescriptorProto

    def msg(name, fields):
        m = f.message_type.add(name=name)
        for fname, num, typ, label, tname, oneof in fields:
            fd = m.field.add(name=fname, number=num, type=typ, label=label)
            if tname:
                fd.type_name = tname
            if oneof is not None:
                fd.oneof_index = oneof
        return m

    msg("BytesList", [("value", 1, T.TYPE_BYTES, T.LABEL_REPEATED, None, None)])
    msg("FloatList", [("value", 1, T.TYPE_FLOAT, T.LABEL_REPEATED, None, None)])
    msg("Int64List", [("value", 1, T.TYPE_INT64, T.LABEL_REPEATED, None, None)])
    feat = msg("Feature", [("bytes_list", 1, T.TYPE_MESSAGE, T.LABEL_OPTIONAL, ".zoo_t.BytesList", 0),
                           ("float_list", 2, T.TYPE_MESSAGE, T.LABEL_OPTIONAL, ".zoo_t.FloatList", 0),
                           ("int64_list", 3, T.TYPE_MESSAGE, T.LABEL_OPTIONAL, ".zoo_t.Int64List", 0)])
    feat.oneof_decl.add(name="kind")
    feats = msg("Features", [("feature", 1, T.TYPE_MESSAGE, T.LABEL_REPEATED, ".zoo_t.Features.FeatureEntry", None)])
    entry = feats.nested_type.add(name="FeatureEntry")
    entry.field.add(name="key", number=1, type=T.TYPE_STRING, label=T.LABEL_OPTIONAL)
    entry.field.add(name="value", number=2, type=T.TYPE_MESSAGE, label=T.LABEL_OPTIONAL, type_name=".zoo_t.Feature")
    entry.options.map_entry = True
    msg("Example", [("features", 1, T.TYPE_MESSAGE, T.LABEL_OPTIONAL, ".zoo_t.Features", None)])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(f)
    Example = message_factory.GetMessageClass(pool.FindMessageTypeByName("zoo_t.Example"))

    ours = encode_example({"x": np.array([1.5, -2.0], np.float32), "n": np.array([3, -7, 2 ** 35], np.int64), "s": [b"ab", b"c"]})
    theirs = Example()
    theirs.ParseFromString(ours)
    assert list(theirs.features.feature["x"].float_list.value) == [1.5, -2.0]
    assert list(theirs.features.feature["n"].int64_list.value) == [3, -7, 2 ** 35]
    assert list(theirs.features.feature["s"].bytes_list.value) == [b"ab", b"c"]
    e = Example()
    e.features.feature["img"].bytes_list.value.append(b"\\xff\\xd8jpeg")
    e.features.feature["y"].int64_list.value.extend([5, -1])
    e.features.feature["w"].float_list.value.extend([0.25, 4.0, 1e-3])
    got = decode_example(e.SerializeToString())
    assert got["img"] == [b"\\xff\\xd8jpeg"] and got["y"].tolist() == [5, -1]
    np.testing.assert_allclose(got["w"], [0.25, 4.0, 1e-3], rtol=1e-6)
