"""TensorFlow frozen-graph import / export without TensorFlow (zoo/tfpark/graphdef.py, TFNet, zoo.util.tf.export_tf / strip_unused).
The GraphDef wire format written here is checked against google.protobuf's own decoder, so files are readable by TensorFlow."""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn as nn


def _cnn():
    torch.manual_seed(0)
    m = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.BatchNorm2d(8), nn.ReLU(), nn.MaxPool2d(2), nn.Conv2d(8, 8, 3, groups=8), nn.Tanh(),
                      nn.AvgPool2d(2), nn.Flatten(), nn.Dropout(0.5), nn.Linear(8 * 3 * 3, 16), nn.LeakyReLU(0.1), nn.Linear(16, 5), nn.Softmax(-1))
    with torch.no_grad():
        m[1].running_mean.uniform_(-0.5, 0.5)
        m[1].running_var.uniform_(0.5, 1.5)
    return m.eval()


def test_export_tf_roundtrip_through_tfnet(tmp_path):
    from zoo.pipeline.api.net import Net
    from zoo.tfpark import TFNet
    from zoo.util.tf import export_tf

    m = _cnn()
    x = torch.randn(4, 3, 16, 16)
    ins, outs = export_tf(m, str(tmp_path / "exp"), x[:1])
    assert ins == ["input:0"] and outs == ["output:0"]
    assert json.load(open(tmp_path / "exp" / "graph_meta.json"))["output_names"] == ["output:0"]
    net = TFNet.from_export_folder(str(tmp_path / "exp"))
    y = net(x.permute(0, 2, 3, 1).contiguous())             # TensorFlow graphs take NHWC
    torch.testing.assert_close(y, m(x), atol=1e-5, rtol=1e-5)
    p = net.predict(x.permute(0, 2, 3, 1).numpy(), batch_per_thread=3)
    np.testing.assert_allclose(p, m(x).detach().numpy(), atol=1e-5)
    g = Net.load_tf(str(tmp_path / "exp"))
    torch.testing.assert_close(g.forward(x.permute(0, 2, 3, 1).contiguous()), m(x), atol=1e-5, rtol=1e-5)
    with pytest.raises(NotImplementedError):
        TFNet.from_session(None, [], [])


def test_graphdef_bytes_are_real_protobuf(tmp_path):
    """Decode the written file with google.protobuf through a dynamically built descriptor of GraphDef's outer layers."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    from zoo.tfpark.graphdef import read_graphdef, write_graphdef

    fd = descriptor_pb2.FileDescriptorProto(name="mini_graph.proto", package="mini", syntax="proto3")
    node = fd.message_type.add(name="NodeDef")
    for i, (n, t, lab) in enumerate([("name", 9, 1), ("op", 9, 1), ("input", 9, 3), ("device", 9, 1)], 1):
        node.field.add(name=n, number=i, type=t, label=lab)
    graph = fd.message_type.add(name="GraphDef")
    graph.field.add(name="node", number=1, type=11, label=3, type_name=".mini.NodeDef")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    GraphDef = message_factory.GetMessageClass(pool.FindMessageTypeByName("mini.GraphDef"))
    data = write_graphdef(str(tmp_path / "g.pb"), [("x", "Placeholder", [], {"dtype": np.float32}),
                                                  ("w", "Const", [], {"dtype": np.float32, "value": np.eye(3, dtype=np.float32)}),
                                                  ("y", "MatMul", ["x", "w"], {"T": np.float32, "transpose_b": True})])
    g = GraphDef()
    g.ParseFromString(data)
    assert [(n.name, n.op, list(n.input)) for n in g.node] == [("x", "Placeholder", []), ("w", "Const", []), ("y", "MatMul", ["x", "w"])]
    nodes = read_graphdef(str(tmp_path / "g.pb"))
    assert nodes[2].a("transpose_b") is True and nodes[1].a("value").shape == (3, 3)


def test_tfnet_ops_and_strip_unused(tmp_path):
    from zoo.tfpark import TFNet
    from zoo.tfpark.graphdef import write_graphdef
    from zoo.util.tf import strip_unused

    f32, i32 = np.float32, np.int32
    rng = np.random.RandomState(0)
    W = rng.randn(6, 4).astype(f32)
    nodes = [("x", "Placeholder", [], {"dtype": f32}), ("unused", "Placeholder", [], {"dtype": f32}),
             ("w", "Const", [], {"dtype": f32, "value": W}), ("axis", "Const", [], {"dtype": i32, "value": np.array([1], i32)}),
             ("mm", "MatMul", ["x", "w"], {"T": f32}), ("sq", "Square", ["mm"], {"T": f32}),
             ("mean", "Mean", ["sq", "axis"], {"T": f32, "keep_dims": True}), ("rs", "Rsqrt", ["mean"], {"T": f32}),
             ("norm", "Mul", ["mm", "rs"], {"T": f32}), ("one", "Const", [], {"dtype": i32, "value": np.array(1, i32)}),
             ("cat", "ConcatV2", ["norm", "mm", "one"], {"T": f32, "N": 2}), ("dead", "Relu", ["unused"], {"T": f32}),
             ("out", "Sigmoid", ["cat"], {"T": f32})]
    write_graphdef(str(tmp_path / "g.pb"), nodes)
    net = TFNet(str(tmp_path / "g.pb"), ["x"], ["out"])
    x = torch.randn(5, 6)
    mm = x @ torch.from_numpy(W)
    want = torch.sigmoid(torch.cat([mm * torch.rsqrt((mm * mm).mean(1, keepdim=True)), mm], 1))
    torch.testing.assert_close(net(x), want, atol=1e-5, rtol=1e-5)
    pruned = strip_unused(str(tmp_path / "g.pb"), ["mm"], ["out"])
    (tmp_path / "p.pb").write_bytes(pruned)
    net2 = TFNet(str(tmp_path / "p.pb"), ["mm"], ["out"])
    assert set(net2.graph.nodes) == {"mm", "axis", "sq", "mean", "rs", "norm", "one", "cat", "out"}
    torch.testing.assert_close(net2(mm), want, atol=1e-5, rtol=1e-5)
    with pytest.raises(NotImplementedError):
        write_graphdef(str(tmp_path / "bad.pb"), [("x", "Placeholder", [], {"dtype": f32}), ("y", "SomeExoticOp", ["x"], {})])
        TFNet(str(tmp_path / "bad.pb"), ["x"], ["y"])(x)


@pytest.mark.skipif(not os.path.exists("/root/reference/zoo/src/test/resources/tfnet/frozen_inference_graph.pb"),
                    reason="reference test resources are not on this machine")
def test_reference_frozen_graph_written_by_tensorflow():
    """The reference's own TFNet fixture -- a graph frozen by real TensorFlow (TFNetSpec.scala loads the same folder) -- through
    the GraphDef reader and `TFNet.from_export_folder`, checked against the same arithmetic done with the graph's constants."""
    import torch

    from zoo.pipeline.api.net import TFNet
    from zoo.tfpark.graphdef import read_graphdef

    folder = "/root/reference/zoo/src/test/resources/tfnet"
    nodes = {n.name: n for n in read_graphdef(folder + "/frozen_inference_graph.pb")}
    assert nodes["dense/MatMul"].op == "MatMul" and nodes["dense_1/Sigmoid"].inputs == ["dense_1/BiasAdd"]
    w1, b1 = nodes["dense/kernel"].a("value"), nodes["dense/bias"].a("value")
    w2, b2 = nodes["dense_1/kernel"].a("value"), nodes["dense_1/bias"].a("value")
    net = TFNet.from_export_folder(folder)
    x = torch.rand(5, w1.shape[0])
    want = torch.sigmoid(torch.relu(x @ torch.from_numpy(w1) + torch.from_numpy(b1)) @ torch.from_numpy(w2) + torch.from_numpy(b2))
    torch.testing.assert_close(net(x), want)


_SM = "/root/reference/zoo/src/test/resources/"


@pytest.mark.skipif(not os.path.exists(_SM + "saved-model-signature/saved_model.pb"), reason="reference test resources are not on this machine")
def test_reference_saved_models_written_by_tensorflow():
    """SavedModel directories written by real TensorFlow (the fixtures of TFNetSpec.scala:138-176 and pyzoo test_tfnet.py:80-85):
    saved_model.pb is decoded, the variables are read from the TensorBundle (LevelDB-format index + data shard), frozen into the
    graph and the result matches a numpy restatement built from the same variables."""
    from zoo.pipeline.api.net import TFNet
    from zoo.tfpark.saved_model import read_bundle, read_saved_model, read_table

    # -- the signature model: Placeholder [?, 4] -> dense(10); ref variables (VariableV2)
    path = _SM + "saved-model-signature"
    meta = read_saved_model(path)[0]
    assert meta.tags == ["serve"] and meta.signatures == {"serving_default": ({"input": "Placeholder:0"}, {"output": "dense/BiasAdd:0"})}
    keys = [k for k, _ in read_table(path + "/variables/variables.index")]
    assert keys == [b"", b"dense/bias", b"dense/kernel"]                      # header entry first, then the variables in key order
    v = read_bundle(path + "/variables/variables")
    x = torch.ones(4, 4)
    want = x.numpy() @ v["dense/kernel"] + v["dense/bias"]
    for kw in ({}, {"inputs": ["Placeholder:0"]}, {"outputs": ["dense/BiasAdd:0"]}, {"signature": "serving_default", "tag": "serve"}):
        out = TFNet.from_saved_model(path, **kw)(x)
        assert out.shape == (4, 10) and np.allclose(out.numpy(), want, atol=1e-6)
    with pytest.raises(ValueError):
        TFNet.from_saved_model(path, tag="train")

    # -- the Keras model with resource variables: flatten -> dense/relu -> BN -> dense/relu -> BN -> dense/softmax;
    #    BatchNorm picks moving statistics through Switch / Merge on the (false) learning phase
    path = _SM + "saved-model-resource"
    v = read_bundle(path + "/variables/variables")
    assert v["dense/kernel"].shape == (784, 64) and v["Adam/iterations"].shape == ()
    net = TFNet.from_saved_model(path, inputs=["flatten_input:0"], outputs=["dense_2/Softmax:0"])
    x = np.random.RandomState(0).rand(6, 28, 28, 1).astype(np.float32)

    def bn(h, name):
        return (h - v[name + "/moving_mean"]) / np.sqrt(v[name + "/moving_variance"] + 1e-3) * v[name + "/gamma"] + v[name + "/beta"]

    h = bn(np.maximum(x.reshape(6, -1) @ v["dense/kernel"] + v["dense/bias"], 0), "batch_normalization_v1")
    h = bn(np.maximum(h @ v["dense_1/kernel"] + v["dense_1/bias"], 0), "batch_normalization_v1_1")
    logits = h @ v["dense_2/kernel"] + v["dense_2/bias"]
    want = np.exp(logits - logits.max(-1, keepdims=True))
    want /= want.sum(-1, keepdims=True)
    assert np.allclose(net.predict(x), want, atol=1e-5)
    assert np.allclose(TFNet.from_saved_model(path)(torch.from_numpy(x)).numpy(), want, atol=1e-5)      # the signature names the same tensors


def test_table_reader_handles_prefix_compression_and_snappy():
    """The LevelDB-format table reader on a hand-built table: shared key prefixes, two data blocks, one of them snappy-compressed."""
    import struct

    from zoo.tfpark.saved_model import _snappy_uncompress, read_table
    from zoo.util.protowire import write_varint as vi

    def block(pairs):
        out, prev = b"", b""
        for k, val in pairs:
            shared = 0
            while shared < min(len(prev), len(k)) and prev[shared] == k[shared]:
                shared += 1
            out += vi(shared) + vi(len(k) - shared) + vi(len(val)) + k[shared:] + val
            prev = k
        return out + struct.pack("<II", 0, 1)                               # one restart point at offset 0

    def snappy_literal(data):                                               # a valid snappy stream made of one literal
        assert len(data) < 60
        return vi(len(data)) + bytes([(len(data) - 1) << 2]) + data

    assert _snappy_uncompress(vi(8) + bytes([(4 - 1) << 2]) + b"abcd" + bytes([(0 << 5) | ((4 - 4) << 2) | 1, 4])) == b"abcdabcd"   # literal + copy
    b1 = block([(b"layer/bias", b"B"), (b"layer/kernel", b"K")])
    b2 = block([(b"other", b"O")])
    file = b1 + b"\x00" + b"\x00" * 4
    off2 = len(file)
    c2 = snappy_literal(b2)
    file += c2 + b"\x01" + b"\x00" * 4
    index = block([(b"layer/kernel", vi(0) + vi(len(b1))), (b"other", vi(off2) + vi(len(c2)))])
    ioff = len(file)
    file += index + b"\x00" + b"\x00" * 4
    footer = vi(0) + vi(0) + vi(ioff) + vi(len(index))
    file += footer + b"\x00" * (40 - len(footer)) + struct.pack("<Q", 0xDB4775248B80FB57)
    import tempfile

    with tempfile.NamedTemporaryFile(suffix=".index", delete=False) as f:
        f.write(file)
    try:
        assert read_table(f.name) == [(b"layer/bias", b"B"), (b"layer/kernel", b"K"), (b"other", b"O")]
    finally:
        os.unlink(f.name)


@pytest.mark.skipif(not os.path.exists(_SM + "saved-model-signature/saved_model.pb"), reason="reference test resources are not on this machine")
def test_saved_model_through_inference_model_and_net_load():
    from zoo.pipeline.api.net import Net
    from zoo.pipeline.inference import InferenceModel

    x = np.ones((3, 4), np.float32)
    a = InferenceModel().load_tensorflow(_SM + "saved-model-signature", model_type="savedModel").predict(x)
    b = InferenceModel().load_tf(_SM + "saved-model-signature").predict(x)                       # recognised by its saved_model.pb
    c = Net.load_tf(_SM + "saved-model-signature")(torch.from_numpy(x)).detach().numpy()
    assert np.asarray(a).shape == (3, 10) and np.allclose(a, b) and np.allclose(a, c, atol=1e-6)


@pytest.mark.skipif(not os.path.exists(_SM + "tfnet_training/graph_meta.json"), reason="reference test resources are not on this machine")
def test_reference_training_export_is_trainable_and_its_gradient_nodes_agree_with_autograd():
    """tfnet_training is an `export_tf(generate_backward=True)` folder written by TensorFlow (TFNetSpec.scala:126-136): forward
    graph + explicit gradient nodes + the list of variables.  The variables become parameters of the imported module; torch autograd
    through the imported forward ops must equal what the graph's own SigmoidGrad / ReluGrad / BiasAddGrad / MatMul nodes compute."""
    import json

    from zoo.pipeline.api.net import TFNet
    from zoo.tfpark.graphdef import GraphDefModule, read_graphdef

    folder = _SM + "tfnet_training"
    meta = json.load(open(folder + "/graph_meta.json"))
    net = TFNet.from_export_folder(folder)
    params = dict(net.named_parameters())
    assert len(params) == 4 and sorted(tuple(p.shape) for p in params.values()) == [(2,), (4, 10), (10,), (10, 2)]
    torch.manual_seed(0)
    x, gy = torch.rand(2, 4, requires_grad=True), torch.rand(2, 2)
    out = net(x)
    out.backward(gy)
    nodes = read_graphdef(folder + "/frozen_inference_graph.pb")
    grads = GraphDefModule(nodes, ["Placeholder:0", "dense_1/Sigmoid_grad:0"], meta["grad_inputs"] + meta["grad_variables"])
    with torch.no_grad():
        gx, *gvars = grads(x.detach(), gy)
    assert torch.allclose(x.grad, gx, atol=1e-6)
    by_name = {net.graph.nodes[n].name: getattr(net.graph, k) for n, k in net.graph._const.items()}
    for var, g in zip(meta["variables"], gvars):
        assert torch.allclose(by_name[var.split(":")[0]].grad, g, atol=1e-6), var
    # and it trains: a few SGD steps on the parameters lower a loss
    opt = torch.optim.SGD(net.parameters(), lr=0.5)
    tgt = torch.tensor([[1.0, 0.0], [0.0, 1.0]])
    first = None
    for _ in range(50):
        opt.zero_grad()
        loss = ((net(x.detach()) - tgt) ** 2).mean()
        loss.backward()
        opt.step()
        first = first if first is not None else float(loss.detach())
    assert float(loss.detach()) < first * 0.7


@pytest.mark.skipif(not os.path.exists(_SM + "tf/multi_type_inputs_outputs.pb"), reason="reference test resources are not on this machine")
def test_reference_multi_type_and_string_graphs():
    """tf/multi_type_inputs_outputs.pb (five placeholders of different dtypes, TFNetSpec.scala:100-124) and tfnet_string (a string
    placeholder through StringToNumber)."""
    from zoo.pipeline.api.net import TFNet

    names = ["float", "double", "int", "long", "uint8"]
    net = TFNet.from_frozen_graph(_SM + "tf/multi_type_inputs_outputs.pb", [f"{n}_input:0" for n in names], [f"{n}_output:0" for n in names])
    ins = [torch.ones(3, 1, dtype=t) * 2 for t in (torch.float32, torch.float64, torch.int32, torch.int64, torch.uint8)]
    outs = net(*ins)
    assert [o.dtype for o in outs] == [i.dtype for i in ins] and all(torch.equal(o, i) for o, i in zip(outs, ins))
    snet = TFNet.from_export_folder(_SM + "tfnet_string")
    got = snet(["1.5", b"2", "3e2"])
    assert got.dtype == torch.float32 and got.tolist() == [1.5, 2.0, 300.0]


@pytest.mark.skipif(not os.path.exists(_SM + "saved-model-signature/saved_model.pb"), reason="reference test resources are not on this machine")
def test_tfnet_for_inference_and_graph_runner():
    from zoo.pipeline.api.net import GraphRunner, TFNet, TFNetForInference

    x = torch.ones(2, 4)
    net = TFNetForInference.from_saved_model(_SM + "saved-model-signature")
    assert isinstance(net, TFNetForInference) and isinstance(net, TFNet)
    want = TFNet.from_saved_model(_SM + "saved-model-signature")(x)
    assert torch.equal(net(x), want)
    runner = GraphRunner(_SM + "tfnet/frozen_inference_graph.pb")
    meta = __import__("json").load(open(_SM + "tfnet/graph_meta.json"))
    out = runner.run([torch.rand(3, 4)], meta["input_names"], meta["output_names"])
    assert len(out) == 1 and out[0].shape[0] == 3
    assert runner.run([torch.rand(5, 4)], meta["input_names"], meta["output_names"])[0].shape[0] == 5 and len(runner._modules) == 1


def test_export_tf_training_graph_round_trip(tmp_path):
    """export_tf(generate_backward=True): the folder has the reference's training-export layout (variables, grad_variables, grad_inputs);
    reloaded, the variables are trainable and the written gradient nodes agree with autograd of the original torch model."""
    import json

    from zoo.pipeline.api.net import TFNet
    from zoo.tfpark.graphdef import GraphDefModule, read_graphdef
    from zoo.util.tf import export_tf

    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(4, 6), torch.nn.Tanh(), torch.nn.Linear(6, 3, bias=False), torch.nn.ReLU(), torch.nn.Linear(3, 2),
            torch.nn.Sigmoid())
    ins, outs = export_tf(model, str(tmp_path), example_input=torch.zeros(1, 4), generate_backward=True)
    meta = json.load(open(tmp_path / "graph_meta.json"))
    assert ins == ["Placeholder:0"] and outs == ["dense_2/Sigmoid:0"] and meta["variables"] == ["dense/kernel:0", "dense/bias:0", "dense_1/kernel:0",
            "dense_2/kernel:0", "dense_2/bias:0"]
    x, gy = torch.randn(5, 4, requires_grad=True), torch.randn(5, 2)
    model(x).backward(gy)
    grads = GraphDefModule(read_graphdef(str(tmp_path / "frozen_inference_graph.pb")), ["Placeholder:0", "dense_2/Sigmoid_grad:0"],
            meta["grad_inputs"] + meta["grad_variables"])
    gx, *gv = grads(x.detach(), gy)
    assert torch.allclose(gx, x.grad, atol=1e-5)
    want = [model[0].weight.grad.t(), model[0].bias.grad, model[2].weight.grad.t(), model[4].weight.grad.t(), model[4].bias.grad]
    assert all(torch.allclose(a, b, atol=1e-5) for a, b in zip(gv, want))
    net = TFNet.from_export_folder(str(tmp_path))
    assert len(list(net.parameters())) == 5 and torch.allclose(net(x.detach()), model(x.detach()), atol=1e-6)
    with pytest.raises(NotImplementedError):
        export_tf(torch.nn.Sequential(torch.nn.Conv2d(1, 1, 3)), str(tmp_path / "c"), example_input=torch.zeros(1, 1, 8, 8), generate_backward=True)
