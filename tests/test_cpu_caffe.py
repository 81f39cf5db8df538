"""Caffe import (zoo/models/caffe): a prototxt + binary caffemodel written here must load into a torch graph that computes what an
equivalent hand-built torch network computes.  Reference behaviour: models/caffe/CaffeLoader.scala, Net.loadCaffe."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

PROTOTXT = '''
name: "MiniNet"
input: "data"
input_dim: 1 input_dim: 3 input_dim: 16 input_dim: 16
layer { name: "conv1" type: "Convolution" bottom: "data" top: "conv1"
        convolution_param { num_output: 8 kernel_size: 3 stride: 1 pad: 1 } }
layer { name: "bn1" type: "BatchNorm" bottom: "conv1" top: "conv1" batch_norm_param { use_global_stats: true eps: 0.001 } }
layer { name: "scale1" type: "Scale" bottom: "conv1" top: "conv1" scale_param { bias_term: true } }
layer { name: "relu1" type: "ReLU" bottom: "conv1" top: "conv1" }
layer { name: "pool1" type: "Pooling" bottom: "conv1" top: "pool1" pooling_param { pool: MAX kernel_size: 3 stride: 2 } }
layer { name: "branch_a" type: "Convolution" bottom: "pool1" top: "a" convolution_param { num_output: 4 kernel_size: 1 } }
layer { name: "branch_b" type: "Convolution" bottom: "pool1" top: "b" convolution_param { num_output: 4 kernel_h: 3 kernel_w: 3 pad_h: 1 pad_w: 1 group: 2 } }
layer { name: "cat" type: "Concat" bottom: "a" bottom: "b" top: "cat" }
layer { name: "sum" type: "Eltwise" bottom: "cat" bottom: "pool1" top: "sum" eltwise_param { operation: SUM coeff: 1 coeff: 0.5 } }
layer { name: "drop" type: "Dropout" bottom: "sum" top: "sum" dropout_param { dropout_ratio: 0.3 } }
layer { name: "fc" type: "InnerProduct" bottom: "sum" top: "fc" inner_product_param { num_output: 5 } }
layer { name: "prob" type: "Softmax" bottom: "fc" top: "prob" }
layer { name: "loss" type: "SoftmaxWithLoss" bottom: "fc" bottom: "label" top: "loss" include { phase: TRAIN } }
'''


def _weights():
    g = torch.Generator().manual_seed(0)
    r = lambda *s: torch.randn(*s, generator=g) * 0.3
    return {"conv1": [r(8, 3, 3, 3), r(8)], "bn1": [r(8), r(8).abs() + 0.5, torch.tensor([2.0])], "scale1": [r(8) + 1, r(8)],
            "branch_a": [r(4, 8, 1, 1), r(4)], "branch_b": [r(4, 4, 3, 3), r(4)], "fc": [r(5, 8 * 8 * 8), r(5)]}


def _expected(x, w):
    y = F.conv2d(x, w["conv1"][0], w["conv1"][1], padding=1)
    mean, var = w["bn1"][0] / 2.0, w["bn1"][1] / 2.0
    y = (y - mean.view(1, -1, 1, 1)) / torch.sqrt(var.view(1, -1, 1, 1) + 1e-3)
    y = F.relu(y * w["scale1"][0].view(1, -1, 1, 1) + w["scale1"][1].view(1, -1, 1, 1))
    p = F.max_pool2d(y, 3, 2, ceil_mode=True)
    a = F.conv2d(p, w["branch_a"][0], w["branch_a"][1])
    b = F.conv2d(p, w["branch_b"][0], w["branch_b"][1], padding=1, groups=2)
    s = torch.cat([a, b], 1) + 0.5 * p
    return torch.softmax(F.linear(s.flatten(1), w["fc"][0], w["fc"][1]), 1)


def test_caffe_prototxt_and_caffemodel_load(tmp_path):
    from zoo.models.caffe import CaffeConversionException, CaffeLoader, load_caffe, save_caffemodel
    from zoo.pipeline.api.net import Net

    w = _weights()
    (tmp_path / "net.prototxt").write_text(PROTOTXT)
    types = {"conv1": "Convolution", "bn1": "BatchNorm", "scale1": "Scale", "branch_a": "Convolution", "branch_b": "Convolution", "fc": "InnerProduct"}
    save_caffemodel(str(tmp_path / "net.caffemodel"), [(n, types[n], bl) for n, bl in w.items()], name="MiniNet")
    net = load_caffe(str(tmp_path / "net.prototxt"), str(tmp_path / "net.caffemodel")).eval()
    assert net.input_names == ["data"] and net.output_names == ["prob"] and not net.unused_weights
    x = torch.randn(2, 3, 16, 16)
    torch.testing.assert_close(net(x), _expected(x, w), atol=1e-5, rtol=1e-5)
    # the public entry point of the reference
    net2 = Net.load_caffe(str(tmp_path / "net.prototxt"), str(tmp_path / "net.caffemodel"))
    torch.testing.assert_close(net2.forward(x), _expected(x, w), atol=1e-5, rtol=1e-5)
    # match_all: a layer with parameters and no weights is an error
    del w["branch_b"]
    save_caffemodel(str(tmp_path / "partial.caffemodel"), [(n, types[n], bl) for n, bl in w.items()])
    with pytest.raises(CaffeConversionException):
        load_caffe(str(tmp_path / "net.prototxt"), str(tmp_path / "partial.caffemodel"))
    # trainable afterwards
    net.train()
    net(x).sum().backward()
    assert all(p.grad is not None for p in net.parameters())
    # copy weights into an existing torch model by module name
    m = nn.ModuleDict({"conv1": nn.Conv2d(3, 8, 3, padding=1)})
    CaffeLoader.load(m, str(tmp_path / "net.prototxt"), str(tmp_path / "net.caffemodel"), match_all=False)
    torch.testing.assert_close(m["conv1"].weight.detach(), _weights()["conv1"][0])


def test_caffe_v1_layers_and_custom_converter(tmp_path):
    """V1 definitions (`layers { type: CONVOLUTION }`) with weights in V1 form, a data layer as input, and a user converter."""
    from zoo.models.caffe import load_caffe
    from zoo.models.caffe import caffe_schema as cs
    from zoo.util import protowire as pw

    proto = '''
name: "V1Net"
layers { name: "data" type: DATA top: "data" top: "label" }
layers { name: "ip1" type: INNER_PRODUCT bottom: "data" top: "ip1" inner_product_param { num_output: 6 } }
layers { name: "act" type: TANH bottom: "ip1" top: "ip1" }
layers { name: "mine" type: MYLAYER bottom: "ip1" top: "out" }
layers { name: "acc" type: ACCURACY bottom: "out" bottom: "label" top: "acc" }
'''
    (tmp_path / "v1.prototxt").write_text(proto)
    g = torch.Generator().manual_seed(1)
    W, b = torch.randn(6, 10, generator=g), torch.randn(6, generator=g)
    msg = {"name": "V1Net", "layers": [{"name": "ip1", "type": 14, "blobs": [
        {"num": 1, "channels": 1, "height": 6, "width": 10, "data": W.reshape(-1).numpy()},
        {"num": 1, "channels": 1, "height": 1, "width": 6, "data": b.numpy()}]}]}
    (tmp_path / "v1.caffemodel").write_bytes(pw.encode(msg, cs.NET))
    net = load_caffe(str(tmp_path / "v1.prototxt"), str(tmp_path / "v1.caffemodel"),
                     customized_converters={"MYLAYER": lambda layer, blobs: nn.Hardtanh(-0.5, 0.5)})
    x = torch.randn(3, 10)
    torch.testing.assert_close(net(x), torch.clamp(torch.tanh(F.linear(x, W, b)), -0.5, 0.5))
    assert net.input_names == ["data"] and net.output_names == ["out"]
