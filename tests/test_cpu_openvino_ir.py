"""OpenVINO IR importer (zoo/pipeline/inference/openvino_ir.py): hand-written IR files of both format generations against the torch
model they describe (reference: InferenceModel.loadOpenVINO, OpenVinoInferenceSupportive.scala:35-57)."""
import numpy as np
import pytest
import torch
import torch.nn as nn


class _Bin:
    def __init__(self):
        self.buf = b""

    def add(self, arr, dtype=np.float32):
        a = np.ascontiguousarray(np.asarray(arr, dtype=dtype))
        off = len(self.buf)
        self.buf += a.tobytes()
        return off, a.nbytes


def _ports(kind, ids_dims):
    return f"<{kind}>" + "".join(f'<port id="{i}" precision="FP32">' + "".join(f"<dim>{d}</dim>" for d in dims) + "</port>" for i,
            dims in ids_dims) + f"</{kind}>"


def test_ir_v10_cnn(tmp_path):
    torch.manual_seed(0)
    conv, fc = nn.Conv2d(3, 4, 3, stride=2, padding=1), nn.Linear(4 * 4 * 4, 5)
    model = nn.Sequential(conv, nn.ReLU(), nn.MaxPool2d(2, 2, ceil_mode=True), nn.Flatten(), fc, nn.Softmax(-1)).eval()
    b = _Bin()
    w_off, w_sz = b.add(conv.weight.detach().numpy())
    cb_off, cb_sz = b.add(conv.bias.detach().numpy().reshape(1, 4, 1, 1))
    shp_off, shp_sz = b.add([0, -1], np.int64)
    fw_off, fw_sz = b.add(fc.weight.detach().numpy().astype(np.float16), np.float16)        # an FP16 constant, as compressed IRs carry
    fb_off, fb_sz = b.add(fc.bias.detach().numpy().reshape(1, 5))
    const = lambda i, name, et, shape, off, sz: (f'<layer id="{i}" name="{name}" type="Const" version="opset1">'
            f'<data element_type="{et}" offset="{off}" shape="{shape}" size="{sz}"/>'
                                                  + _ports("output", [(0, [])]) + "</layer>")
    layers = [
        f'<layer id="0" name="data" type="Parameter" version="opset1"><data element_type="f32" shape="1,3,16,16"/>'
                f'{_ports("output", [(0, [1, 3, 16, 16])])}</layer>',
        const(1, "conv_w", "f32", "4,3,3,3", w_off, w_sz),
        f'<layer id="2" name="conv" type="Convolution" version="opset1">'
        f'<data auto_pad="explicit" dilations="1,1" pads_begin="1,1" pads_end="1,1" strides="2,2"/>'
        f'{_ports("input", [(0, []), (1, [])])}{_ports("output", [(2, [])])}</layer>',
        const(3, "conv_b", "f32", "1,4,1,1", cb_off, cb_sz),
        f'<layer id="4" name="bias" type="Add" version="opset1">{_ports("input", [(0, []), (1, [])])}{_ports("output", [(2, [])])}</layer>',
        f'<layer id="5" name="relu" type="Relu" version="opset1">{_ports("input", [(0, [])])}{_ports("output", [(1, [])])}</layer>',
        f'<layer id="6" name="pool" type="MaxPool" version="opset1">'
        f'<data auto_pad="explicit" kernel="2,2" pads_begin="0,0" pads_end="0,0" rounding_type="ceil" strides="2,2"/>'
        f'{_ports("input", [(0, [])])}{_ports("output", [(1, [])])}</layer>',
        const(7, "shape", "i64", "2", shp_off, shp_sz),
        f'<layer id="8" name="flatten" type="Reshape" version="opset1"><data special_zero="true"/>{_ports("input", [(0, []), (1, [])])}'
                f'{_ports("output", [(2, [])])}</layer>',
        const(9, "fc_w", "f16", "5,64", fw_off, fw_sz),
        f'<layer id="10" name="fc" type="MatMul" version="opset1"><data transpose_a="false" transpose_b="true"/>'
                f'{_ports("input", [(0, []), (1, [])])}{_ports("output", [(2, [])])}</layer>',
        const(11, "fc_b", "f32", "1,5", fb_off, fb_sz),
        f'<layer id="12" name="fc_bias" type="Add" version="opset1">{_ports("input", [(0, []), (1, [])])}{_ports("output", [(2, [])])}</layer>',
        f'<layer id="13" name="prob" type="Softmax" version="opset1"><data axis="1"/>{_ports("input", [(0, [])])}{_ports("output", [(1, [])])}</layer>',
        f'<layer id="14" name="result" type="Result" version="opset1">{_ports("input", [(0, [])])}</layer>',
    ]
    edges = [(0, 0, 2, 0), (1, 0, 2, 1), (2, 2, 4, 0), (3, 0, 4, 1), (4, 2, 5, 0), (5, 1, 6, 0), (6, 1, 8, 0), (7, 0, 8, 1), (8, 2, 10, 0), (9, 0, 10, 1),
             (10, 2, 12, 0), (11, 0, 12, 1), (12, 2, 13, 0), (13, 1, 14, 0)]
    xml = ('<?xml version="1.0" ?><net name="cnn" version="10"><layers>' + "".join(layers) + "</layers><edges>" +
           "".join(f'<edge from-layer="{a}" from-port="{p}" to-layer="{c}" to-port="{q}"/>' for a, p, c, q in edges) + "</edges></net>")
    (tmp_path / "model.xml").write_text(xml)
    (tmp_path / "model.bin").write_bytes(b.buf)

    from zoo.pipeline.inference import InferenceModel
    from zoo.pipeline.inference.openvino_ir import load_openvino_ir

    net = load_openvino_ir(str(tmp_path / "model.xml"))
    x = torch.randn(3, 3, 16, 16)
    with torch.no_grad():
        want = model(x)
        got = net(x)
    assert got.shape == (3, 5) and torch.allclose(got, want, atol=2e-3)                      # fc weights went through fp16
    im = InferenceModel().load_openvino(str(tmp_path / "model.xml"), str(tmp_path / "model.bin"))
    assert np.allclose(im.predict(x.numpy()), want.numpy(), atol=2e-3)


def test_ir_v7_blobs_format(tmp_path):
    """The generation of the reference (OpenVINO 2018-2019): weights as <blobs> of Caffe-like layers."""
    torch.manual_seed(1)
    conv, fc = nn.Conv2d(2, 3, 3, padding=1), nn.Linear(3 * 4 * 4, 4)
    scale, shift = torch.rand(3) + 0.5, torch.randn(3)

    def reference(x):
        y = torch.relu(conv(x)) * scale.view(1, 3, 1, 1) + shift.view(1, 3, 1, 1)
        y = nn.functional.avg_pool2d(y, 2, 2)
        return torch.softmax(fc(y.flatten(1)), -1) + 0.0

    b = _Bin()
    cw, cb = b.add(conv.weight.detach().numpy()), b.add(conv.bias.detach().numpy())
    sw, sb = b.add(scale.numpy()), b.add(shift.numpy())
    fw, fb = b.add(fc.weight.detach().numpy()), b.add(fc.bias.detach().numpy())
    blob = lambda w, bb: f'<blobs><weights offset="{w[0]}" size="{w[1]}"/><biases offset="{bb[0]}" size="{bb[1]}"/></blobs>'
    layers = [
        f'<layer id="0" name="data" type="Input" precision="FP32">{_ports("output", [(0, [1, 2, 8, 8])])}</layer>',
        f'<layer id="1" name="conv1" type="Convolution" precision="FP32">'
        f'<data dilations="1,1" group="1" kernel="3,3" output="3" pads_begin="1,1" pads_end="1,1" strides="1,1"/>'
        f'{_ports("input", [(0, [])])}{_ports("output", [(1, [])])}{blob(cw, cb)}</layer>',
        f'<layer id="2" name="relu1" type="ReLU" precision="FP32">{_ports("input", [(0, [])])}{_ports("output", [(1, [])])}</layer>',
        f'<layer id="3" name="scale1" type="ScaleShift" precision="FP32">{_ports("input", [(0, [])])}{_ports("output", [(1, [])])}{blob(sw, sb)}</layer>',
        f'<layer id="4" name="pool1" type="Pooling" precision="FP32">'
        f'<data exclude-pad="false" kernel="2,2" pads_begin="0,0" pads_end="0,0" pool-method="avg" rounding_type="floor" strides="2,2"/>'
        f'{_ports("input", [(0, [])])}{_ports("output", [(1, [])])}</layer>',
        f'<layer id="5" name="fc1" type="FullyConnected" precision="FP32"><data out-size="4"/>{_ports("input", [(0, [])])}'
                f'{_ports("output", [(1, [])])}{blob(fw, fb)}</layer>',
        f'<layer id="6" name="prob" type="SoftMax" precision="FP32"><data axis="1"/>{_ports("input", [(0, [])])}{_ports("output", [(1, [])])}</layer>',
    ]
    edges = [(0, 0, 1, 0), (1, 1, 2, 0), (2, 1, 3, 0), (3, 1, 4, 0), (4, 1, 5, 0), (5, 1, 6, 0)]
    xml = ('<?xml version="1.0" ?><net batch="1" name="old" version="6"><layers>' + "".join(layers) + "</layers><edges>" +
           "".join(f'<edge from-layer="{a}" from-port="{p}" to-layer="{c}" to-port="{q}"/>' for a, p, c, q in edges) + "</edges></net>")
    (tmp_path / "old.xml").write_text(xml)
    (tmp_path / "old.bin").write_bytes(b.buf)

    from zoo.pipeline.inference import AbstractInferenceModel
    from zoo.pipeline.inference.openvino_ir import load_openvino_ir

    net = load_openvino_ir(str(tmp_path / "old.xml"))
    x = torch.randn(2, 2, 8, 8)
    with torch.no_grad():
        assert torch.allclose(net(x), reference(x), atol=1e-5)
    m = AbstractInferenceModel()
    m.loadOpenVINO(str(tmp_path / "old.xml"), str(tmp_path / "old.bin"))
    with torch.no_grad():
        assert np.allclose(m.predict(x.numpy()), reference(x).numpy(), atol=1e-5)


def test_ir_details():
    from zoo.pipeline.inference.openvino_ir import _Layer, _fake_quantize, _pool
    import xml.etree.ElementTree as ET

    mk = lambda s: _Layer(ET.fromstring(s))
    x = torch.arange(25.0).view(1, 1, 5, 5)
    same = mk('<layer id="1" name="p" type="MaxPool"><data auto_pad="same_upper" kernel="3,3" strides="2,2" rounding_type="floor"/></layer>')
    assert _pool(same, x, "max").shape == (1, 1, 3, 3)
    ex = mk('<layer id="1" name="p" type="AvgPool"><data kernel="2,2" strides="2,2" pads_begin="1,1" pads_end="0,0" exclude-pad="true"/></layer>')
    got = _pool(ex, torch.ones(1, 1, 4, 4), "avg")
    assert torch.allclose(got, torch.ones_like(got))                                        # padding excluded from the mean
    fq = mk('<layer id="1" name="q" type="FakeQuantize"><data levels="256"/></layer>')
    t = torch.linspace(-2, 2, 9)
    y = _fake_quantize(fq, t, torch.tensor(-1.0), torch.tensor(1.0), torch.tensor(-1.0), torch.tensor(1.0))
    assert float(y.min()) == -1.0 and float(y.max()) == 1.0 and torch.allclose(y[2:7], t[2:7], atol=1 / 255 + 1e-6)
    with pytest.raises(NotImplementedError):
        from zoo.pipeline.inference.openvino_ir import OpenVINOModule

        from types import SimpleNamespace

        OpenVINOModule._new(SimpleNamespace(_linear={}), mk('<layer id="1" name="x" type="NoSuchOp" version="opset1"/>'), [])


def test_ir_linear_layers_are_modules_and_writer_round_trip(tmp_path):
    """Constant-weight matrix products of an IR are nn.Linear children (what `InferenceModel.quantize` replaces by FP8 layers);
    `save_openvino_ir` writes IR v10 that reads back to the same function."""
    from zoo.pipeline.inference.openvino_ir import load_openvino_ir, save_openvino_ir

    torch.manual_seed(0)
    net = nn.Sequential(nn.Conv2d(3, 8, 3, stride=2, padding=1), nn.ReLU(), nn.Conv2d(8, 8, 3, padding=1, groups=4), nn.Tanh(),
                        nn.AvgPool2d(2, 2, padding=1, count_include_pad=False), nn.MaxPool2d(3, 2, ceil_mode=True), nn.Flatten(), nn.Linear(8 * 2 * 2, 10),
                        nn.Sigmoid(), nn.Dropout(0.5), nn.Linear(10, 4, bias=False), nn.Softmax(-1)).eval()
    xml, weights = save_openvino_ir(net, str(tmp_path / "m.xml"), (1, 3, 16, 16))
    ir = load_openvino_ir(xml, weights)
    x = torch.randn(5, 3, 16, 16)
    with torch.no_grad():
        assert torch.allclose(ir(x), net(x), atol=1e-6)
    linears = [m for m in ir.modules() if isinstance(m, nn.Linear)]
    assert sorted((m.in_features, m.out_features) for m in linears) == [(10, 4), (32, 10)]
    with pytest.raises(NotImplementedError):
        save_openvino_ir(nn.Sequential(nn.LSTM(3, 3)), str(tmp_path / "bad.xml"), (1, 3))
