"""TensorFlow graph export helpers (reference: pyzoo/zoo/util/tf.py:50-300: `export_tf` freezes a session graph into a folder
TFNet can load, `strip_unused` prunes a GraphDef to the sub-graph between given inputs and outputs; Scala side
`Net.saveToTf`, pipeline/api/Net.scala).

No TensorFlow runtime exists on this stack, so there is no `tf.Session` to freeze.  What is exported instead is a torch model:
`export_tf(model, folder, example_input)` walks a feed-forward torch module (Linear / Conv2d / pooling / BatchNorm2d /
activations / Flatten / Dropout, nested Sequentials, and this framework's FusedDense) and writes the SAME artefacts the
reference writes -- `frozen_inference_graph.pb` (a real TensorFlow GraphDef, constants frozen in, NHWC convolutions) and
`graph_meta.json` with the input / output tensor names -- so that TensorFlow itself, the reference's TFNet, or this
framework's `TFNet.from_export_folder` can load it.  `strip_unused` works on GraphDef files / bytes directly.
"""
from __future__ import annotations

import json
import os

import numpy as np


def process_grad(grad):
    """The reference replaces `IndexedSlices` / None gradients by dense tensors (tf.py:238-252).  Torch gradients are already
    dense tensors or None: sparse ones are densified, everything else passes through."""
    if grad is None:
        return None
    return grad.to_dense() if getattr(grad, "is_sparse", False) else grad


def _flatten_modules(m):
    import torch.nn as nn

    if isinstance(m, nn.Sequential):
        out = []
        for c in m:
            out += _flatten_modules(c)
        return out
    return [m]


def export_tf(model, folder, example_input=None, inputs=None, outputs=None, generate_backward=False,
              allow_non_differentiable_input=True):
    """Write `folder/frozen_inference_graph.pb` + `folder/graph_meta.json` for a feed-forward torch `model`.

    example_input: a tensor / array with the model's input shape (NCHW for convolutional models; the exported graph takes NHWC,
    as TensorFlow graphs do, and its first Flatten reorders accordingly).  Returns (input_names, output_names)."""
    import torch
    import torch.nn as nn

    from ..tfpark.graphdef import write_graphdef

    if example_input is None or not isinstance(model, nn.Module):
        raise NotImplementedError("export_tf(sess, ...) freezes a live tf.Session, and there is no TensorFlow runtime here; pass a torch "
                                  "model and an example input: export_tf(model, folder, example_input)")
    if generate_backward:
        return _export_training_graph(model, folder)
    x = torch.as_tensor(np.asarray(example_input)) if not torch.is_tensor(example_input) else example_input
    model = model.eval()
    nodes = [("input", "Placeholder", [], {"dtype": np.float32})]
    cur, k = "input", 0
    conv_layout = x.dim() == 4          # NHWC inside the graph while the tensor is an image
    shape = list(x.shape)               # torch-side shape (NCHW) of the running activation

    def const(name, arr):
        nodes.append((name, "Const", [], {"dtype": arr.dtype, "value": np.ascontiguousarray(arr)}))
        return name

    def unary(op, attrs=None):
        nonlocal cur, k
        k += 1
        nodes.append((f"{op.lower()}_{k}", op, [cur], dict({"T": np.float32}, **(attrs or {}))))
        cur = f"{op.lower()}_{k}"

    with torch.no_grad():
        probe = x.float()
        for m in _flatten_modules(model):
            name = type(m).__name__
            if hasattr(m, "weight_bf16") and hasattr(m, "in_features"):           # zoo FusedDense
                lin = nn.Linear(m.in_features, m.out_features, bias=m.bias is not None)
                lin.weight.copy_(m.weight.detach().float().cpu())
                if m.bias is not None:
                    lin.bias.copy_(m.bias.detach().float().cpu())
                act = m.activation
                mods = [lin] + ({"relu": [nn.ReLU()], "sigmoid": [nn.Sigmoid()], "tanh": [nn.Tanh()], None: []}.get(act) or [])
            else:
                mods = [m]
            for mm in mods:
                k += 1
                if isinstance(mm, nn.Linear):
                    w = const(f"w_{k}", mm.weight.detach().float().cpu().numpy().T.copy())
                    nodes.append((f"matmul_{k}", "MatMul", [cur, w], {"T": np.float32, "transpose_a": False, "transpose_b": False}))
                    cur = f"matmul_{k}"
                    if mm.bias is not None:
                        b = const(f"b_{k}", mm.bias.detach().float().cpu().numpy())
                        nodes.append((f"biasadd_{k}", "BiasAdd", [cur, b], {"T": np.float32, "data_format": "NHWC"}))
                        cur = f"biasadd_{k}"
                elif isinstance(mm, nn.Conv2d):
                    if mm.padding_mode != "zeros" or isinstance(mm.padding, str):
                        raise NotImplementedError("export_tf: only zero padding given as numbers is supported for Conv2d")
                    ph, pw_ = mm.padding
                    if ph or pw_:
                        p = const(f"pad_{k}", np.array([[0, 0], [ph, ph], [pw_, pw_], [0, 0]], dtype=np.int32))
                        nodes.append((f"padded_{k}", "Pad", [cur, p], {"T": np.float32}))
                        cur = f"padded_{k}"
                    wt = mm.weight.detach().float().cpu().numpy()
                    if mm.groups == 1:
                        w = const(f"w_{k}", wt.transpose(2, 3, 1, 0).copy())
                        op = "Conv2D"
                    elif mm.groups == mm.in_channels:
                        mult = mm.out_channels // mm.in_channels
                        w = const(f"w_{k}", wt.reshape(mm.in_channels, mult, *wt.shape[2:]).transpose(2, 3, 0, 1).copy())
                        op = "DepthwiseConv2dNative"
                    else:
                        raise NotImplementedError("export_tf: grouped convolutions other than depthwise are not expressible in a TF graph")
                    nodes.append((f"conv_{k}", op, [cur, w], {"T": np.float32, "strides": [1, mm.stride[0], mm.stride[1], 1], "padding": "VALID",
                                                                "data_format": "NHWC", "dilations": [1, mm.dilation[0], mm.dilation[1], 1]}))
                    cur = f"conv_{k}"
                    if mm.bias is not None:
                        b = const(f"b_{k}", mm.bias.detach().float().cpu().numpy())
                        nodes.append((f"biasadd_{k}", "BiasAdd", [cur, b], {"T": np.float32, "data_format": "NHWC"}))
                        cur = f"biasadd_{k}"
                elif isinstance(mm, (nn.MaxPool2d, nn.AvgPool2d)):
                    ks = mm.kernel_size if isinstance(mm.kernel_size, tuple) else (mm.kernel_size, mm.kernel_size)
                    st = mm.stride if isinstance(mm.stride, tuple) else (mm.stride or ks[0], mm.stride or ks[1])
                    pd = mm.padding if isinstance(mm.padding, tuple) else (mm.padding, mm.padding)
                    if any(pd) or getattr(mm, "ceil_mode", False):
                        raise NotImplementedError("export_tf: pooling with padding / ceil_mode is not supported")
                    nodes.append((f"pool_{k}", "MaxPool" if isinstance(mm, nn.MaxPool2d) else "AvgPool", [cur],
                                  {"T": np.float32, "ksize": [1, ks[0], ks[1], 1], "strides": [1, st[0], st[1], 1], "padding": "VALID", "data_format": "NHWC"}))
                    cur = f"pool_{k}"
                elif isinstance(mm, nn.BatchNorm2d):
                    c = mm.num_features
                    g = mm.weight.detach().float().cpu().numpy() if mm.affine else np.ones(c, np.float32)
                    be = mm.bias.detach().float().cpu().numpy() if mm.affine else np.zeros(c, np.float32)
                    ins = [cur, const(f"g_{k}", g), const(f"be_{k}", be), const(f"mu_{k}", mm.running_mean.float().cpu().numpy()),
                           const(f"var_{k}", mm.running_var.float().cpu().numpy())]
                    nodes.append((f"bn_{k}", "FusedBatchNormV3", ins, {"T": np.float32, "U": np.float32, "epsilon": float(mm.eps),
                            "data_format": "NHWC", "is_training": False}))
                    cur = f"bn_{k}"
                elif isinstance(mm, nn.Flatten):
                    if conv_layout and len(shape) == 4:        # the torch model flattens NCHW: put the graph's NHWC tensor in that order first
                        p = const(f"perm_{k}", np.array([0, 3, 1, 2], dtype=np.int32))
                        nodes.append((f"nchw_{k}", "Transpose", [cur, p], {"T": np.float32, "Tperm": np.int32}))
                        cur = f"nchw_{k}"
                    s = const(f"shape_{k}", np.array([-1, int(np.prod(shape[1:]))], dtype=np.int32))
                    nodes.append((f"flatten_{k}", "Reshape", [cur, s], {"T": np.float32, "Tshape": np.int32}))
                    cur = f"flatten_{k}"
                    conv_layout = False
                elif isinstance(mm, (nn.ReLU, nn.ReLU6, nn.Sigmoid, nn.Tanh, nn.ELU, nn.Softplus)):
                    unary({"ReLU": "Relu", "ReLU6": "Relu6", "Sigmoid": "Sigmoid", "Tanh": "Tanh", "ELU": "Elu", "Softplus": "Softplus"}[type(mm).__name__])
                elif isinstance(mm, nn.LeakyReLU):
                    unary("LeakyRelu", {"alpha": float(mm.negative_slope)})
                elif isinstance(mm, nn.Softmax):
                    unary("Softmax")
                elif isinstance(mm, nn.LogSoftmax):
                    unary("LogSoftmax")
                elif isinstance(mm, (nn.Dropout, nn.Dropout2d, nn.Identity)):
                    pass
                else:
                    raise NotImplementedError(f"export_tf: no TensorFlow counterpart registered for layer {name}")
                probe = mm(probe)
                shape = list(probe.shape)
    nodes.append(("output", "Identity", [cur], {"T": np.float32}))
    os.makedirs(folder, exist_ok=True)
    write_graphdef(os.path.join(folder, "frozen_inference_graph.pb"), nodes)
    meta = {"input_names": ["input:0"], "output_names": ["output:0"]}
    with open(os.path.join(folder, "graph_meta.json"), "w") as f:
        json.dump(meta, f)
    return meta["input_names"], meta["output_names"]


def _export_training_graph(model, folder):
    """`export_tf(..., generate_backward=True)`: forward graph + explicit gradient nodes + the metadata a trainable TFNet needs
    (`variables`, `grad_variables`, `grad_inputs`, `temp_tensors`), in the layout of the reference's training exports
    (pyzoo/zoo/util/tf.py:120-238; fixture zoo/src/test/resources/tfnet_training).  Covered: chains of Linear / ReLU / Sigmoid / Tanh --
    the gradient of anything else would need TensorFlow's gradient registry."""
    import torch.nn as nn

    from ..tfpark.graphdef import write_graphdef

    f32 = np.float32
    layers = _flatten_modules(model.eval())
    nodes, variables, temp = [("Placeholder", "Placeholder", [], {"dtype": f32})], [], []
    cur, k, tape = "Placeholder", 0, []                       # tape: what the backward pass needs, in forward order
    for m in layers:
        if isinstance(m, nn.Linear):
            name = "dense" if k == 0 else f"dense_{k}"
            k += 1
            nodes += [(f"{name}/kernel", "Const", [], {"dtype": f32, "value": np.ascontiguousarray(m.weight.detach().float().cpu().numpy().T)}),
                      (f"{name}/kernel/read", "Identity", [f"{name}/kernel"], {"T": f32}),
                      (f"{name}/MatMul", "MatMul", [cur, f"{name}/kernel/read"], {"T": f32, "transpose_a": False, "transpose_b": False})]
            variables.append(f"{name}/kernel:0")
            temp.append(f"{name}/kernel/read:0")
            tape.append(("matmul", name, cur))
            cur = f"{name}/MatMul"
            if m.bias is not None:
                nodes += [(f"{name}/bias", "Const", [], {"dtype": f32, "value": m.bias.detach().float().cpu().numpy().copy()}),
                          (f"{name}/bias/read", "Identity", [f"{name}/bias"], {"T": f32}),
                          (f"{name}/BiasAdd", "BiasAdd", [cur, f"{name}/bias/read"], {"T": f32, "data_format": "NHWC"})]
                variables.append(f"{name}/bias:0")
                tape.append(("bias", name, cur))
                cur = f"{name}/BiasAdd"
        elif isinstance(m, (nn.ReLU, nn.Sigmoid, nn.Tanh)):
            op = {"ReLU": "Relu", "Sigmoid": "Sigmoid", "Tanh": "Tanh"}[type(m).__name__]
            last = "dense" if k <= 1 else f"dense_{k - 1}"
            nodes.append((f"{last}/{op}", op, [cur], {"T": f32}))
            tape.append(("act", op, f"{last}/{op}"))
            cur = f"{last}/{op}"
            temp.append(cur + ":0")
        elif isinstance(m, (nn.Dropout, nn.Identity)):
            continue
        else:
            raise NotImplementedError(f"export_tf(generate_backward=True): no gradient written for layer {type(m).__name__} "
                                      "(covered: Linear, ReLU, Sigmoid, Tanh)")
    out = cur
    nodes += [(f"{out}_grad", "Placeholder", [], {"dtype": f32}), ("gradients/grad_ys_0", "Identity", [f"{out}_grad"], {"T": f32})]
    g, grad_of = "gradients/grad_ys_0", {}
    for kind, a, b in reversed(tape):
        if kind == "act":
            gop = {"Relu": "ReluGrad", "Sigmoid": "SigmoidGrad", "Tanh": "TanhGrad"}[a]
            name = f"gradients/{b}_grad/{gop}"
            nodes.append((name, gop, [g, b] if a == "Relu" else [b, g], {"T": f32}))       # ReluGrad(gradients, features); Sigmoid/TanhGrad(y, dy)
            g = name
        elif kind == "bias":
            name = f"gradients/{a}/BiasAdd_grad/BiasAddGrad"
            nodes.append((name, "BiasAddGrad", [g], {"T": f32, "data_format": "NHWC"}))
            grad_of[f"{a}/bias:0"] = name + ":0"
        else:
            dx, dw = f"gradients/{a}/MatMul_grad/MatMul", f"gradients/{a}/MatMul_grad/MatMul_1"
            nodes += [(dx, "MatMul", [g, f"{a}/kernel/read"], {"T": f32, "transpose_a": False, "transpose_b": True}),
                      (dw, "MatMul", [b, g], {"T": f32, "transpose_a": True, "transpose_b": False})]
            grad_of[f"{a}/kernel:0"] = dw + ":0"
            g = dx
    os.makedirs(folder, exist_ok=True)
    write_graphdef(os.path.join(folder, "frozen_inference_graph.pb"), nodes)
    meta = {"input_names": ["Placeholder:0"], "output_names": [out + ":0"], "variables": variables, "grad_variables": [grad_of[v] for v in variables],
            "grad_inputs": [g + ":0"], "temp_tensors": temp}
    with open(os.path.join(folder, "graph_meta.json"), "w") as f:
        json.dump(meta, f)
    return meta["input_names"], meta["output_names"]


def strip_unused(input_graph_def, input_tensor_names, output_tensor_names, placeholder_type_enum=1):
    """Prune a GraphDef (path or bytes) to what `output_tensor_names` need, cutting at `input_tensor_names`, which become
    Placeholders (reference: tf.py strip_unused, after TensorFlow's strip_unused_lib).  Returns the pruned GraphDef bytes."""
    from ..tfpark.graphdef import DT, read_graphdef, write_graphdef

    nodes = {n.name: n for n in read_graphdef(input_graph_def)}
    strip = lambda s: s.split(":")[0].lstrip("^")
    cut = {strip(n) for n in input_tensor_names}
    keep, stack = [], [strip(n) for n in output_tensor_names]
    seen = set()
    while stack:
        name = stack.pop()
        if name in seen:
            continue
        seen.add(name)
        if name not in nodes:
            raise KeyError(f"{name} is not in the graph")
        keep.append(name)
        if name not in cut:
            stack += [strip(i) for i in nodes[name].inputs]
    dtype = DT.get(placeholder_type_enum, np.float32)
    out = []
    for n in read_graphdef(input_graph_def):          # original order
        if n.name not in seen:
            continue
        out.append((n.name, "Placeholder", [], {"dtype": dtype}) if n.name in cut else node_tuple(n))
    return write_graphdef(None, out)


_TYPE_ATTRS = ("T", "dtype", "DstT", "SrcT", "Tshape", "Tperm", "U", "Tidx", "Tpaddings", "Tparams", "Tindices", "Taxis")


def node_tuple(n):
    """A parsed `graphdef.Node` -> the (name, op, inputs, attrs) tuple `write_graphdef` takes."""
    from ..tfpark.graphdef import DT

    attrs = {}
    for key in n.attr:
        v = n.a(key)
        if isinstance(v, int) and key in _TYPE_ATTRS:
            v = DT.get(v, np.float32)
        if v is not None and not (isinstance(v, list) and len(v) == 0):
            attrs[key] = v
    return (n.name, n.op, list(n.inputs), attrs)
