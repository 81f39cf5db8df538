"""FP8 inference path, CPU oracle side: quantised layers reproduce the fp32 model within FP8 resolution, chains pass E4M3 on."""
import copy

import torch
import torch.nn as nn


def test_quantize_linear_layers_cpu_oracle():
    from zoo.ops.dense import FusedMLP
    from zoo.ops.quant import FP8Linear, FP8Tensor, dequantize_e4m3, quantize_e4m3, quantize_linear_layers

    torch.manual_seed(0)
    x = torch.randn(64, 40)
    q = quantize_e4m3(x, float(x.abs().max()) / 448.0)
    assert q.shape == (64, 48) and q.dtype == torch.uint8 and int(q[:, 40:].sum()) == 0
    back = dequantize_e4m3(q, float(x.abs().max()) / 448.0, 40)
    assert float((back - x).abs().max()) < 0.07 * float(x.abs().max())          # 3 mantissa bits
    model = nn.Sequential(FusedMLP(40, [64, 32], activation="relu"), nn.Linear(32, 5), nn.Softmax(-1))
    ref = copy.deepcopy(model).eval()
    qm = quantize_linear_layers(model, x)
    kinds = [type(m).__name__ for m in qm.modules()]
    assert kinds.count("FP8Linear") == 3 and "FusedDense" not in kinds
    assert qm[0][0].next_scale == qm[0][1].in_scale and qm[0][1].next_scale is None      # the tower is chained, its end is not
    assert isinstance(qm[0][0](x), FP8Tensor)
    got, want = qm(x), ref(x)
    assert got.shape == want.shape and float((got - want).abs().max()) < 0.05
    assert (got.argmax(1) == want.argmax(1)).float().mean() > 0.9
    assert isinstance(qm[1], FP8Linear) and qm[1].wq.shape == (5, 32)


def test_inference_model_quantize_and_foreign_loaders(tmp_path):
    """InferenceModel.quantize (the int8 route of the reference) and the caffe / tensorflow loaders, on CPU."""
    import numpy as np

    from zoo.models.caffe import save_caffemodel
    from zoo.pipeline.inference import InferenceModel
    from zoo.util.tf import export_tf

    torch.manual_seed(0)
    net = nn.Sequential(nn.Linear(12, 32), nn.ReLU(), nn.Linear(32, 4), nn.Softmax(-1)).eval()
    x = torch.randn(50, 12)
    im = InferenceModel(device="cpu").load_torch(copy.deepcopy(net))
    want = im.predict(x.numpy())
    im.quantize(x)
    got = im.predict(x.numpy())
    assert np.abs(got - want).max() < 0.05 and sum(type(m).__name__ == "FP8Linear" for m in im.model.modules()) == 2
    # tensorflow frozen graph
    export_tf(net, str(tmp_path / "tf"), x[:1])
    tfm = InferenceModel(device="cpu").load_tf(str(tmp_path / "tf"))
    np.testing.assert_allclose(tfm.predict(x.numpy()), want, atol=1e-5)
    # caffe
    (tmp_path / "n.prototxt").write_text('name: "n" input: "data" input_dim: 1 input_dim: 12\n'
                                         'layer { name: "ip" type: "InnerProduct" bottom: "data" top: "ip" inner_product_param { num_output: 32 } }\n'
                                         'layer { name: "r" type: "ReLU" bottom: "ip" top: "ip" }\n'
                                         'layer { name: "ip2" type: "InnerProduct" bottom: "ip" top: "ip2" inner_product_param { num_output: 4 } }\n'
                                         'layer { name: "p" type: "Softmax" bottom: "ip2" top: "p" }\n')
    save_caffemodel(str(tmp_path / "n.caffemodel"), [("ip", "InnerProduct", [net[0].weight, net[0].bias]), ("ip2", "InnerProduct", [net[2].weight,
            net[2].bias])])
    cm = InferenceModel(device="cpu").load_caffe(str(tmp_path / "n.prototxt"), str(tmp_path / "n.caffemodel"))
    np.testing.assert_allclose(cm.predict(x.numpy()), want, atol=1e-5)
