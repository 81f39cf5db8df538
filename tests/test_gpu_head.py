"""K5b: the one-pass classification head (last Dense + softmax + NLL + head backward) against a plain fp32 torch reference, and
the engine step with the fused head against the layer-by-layer step."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,K,C,relu_cols,zero_based,padding", [
    (1000, 128, 5, 64, False, -1),        # the NeuralCF head: [tower | GMF], one-based ratings, default padding value
    (4099, 128, 5, 64, False, None),
    (513, 64, 2, 64, True, None),
    (257, 256, 8, 0, True, 3),
    (31, 40, 3, 16, True, None),
    (70000, 128, 5, 64, False, -1),
])
@pytest.mark.parametrize("impl", [0, 1])          # 0: warp-MMA kernel, 1: SIMT kernel
def test_head_kernel_matches_torch(B, K, C, relu_cols, zero_based, padding, impl):
    from zoo.ops.head_loss import head_reference, head_softmax_nll

    torch.manual_seed(B + K + C)
    dev = "cuda"
    x = torch.randn(B, K, device=dev)
    x[:, :relu_cols] = x[:, :relu_cols].relu()
    x16 = x.to(torch.bfloat16)
    w = (torch.randn(C, K, device=dev) * 0.2)
    w16 = torch.zeros(C, (K + 7) // 8 * 8, dtype=torch.bfloat16, device=dev)
    w16[:, :K] = w
    bias = torch.randn(C, device=dev) * 0.1
    labels = torch.randint(0, C, (B,), device=dev) + (0 if zero_based else 1)
    if padding is not None:
        labels[::7] = padding
    dw = torch.zeros(C, K, device=dev)
    db = torch.zeros(C, device=dev)
    out, dx, logits = head_softmax_nll(x16, w16, bias, labels, dw, db, zero_based, padding, True, relu_cols, want_logits=True, impl=impl)
    loss_r, dx_r, dw_r, db_r, logits_r = head_reference(x16, w16[:, :K], bias, labels, zero_based, padding, True, relu_cols)
    torch.testing.assert_close(logits, logits_r, atol=2e-3, rtol=1e-3)
    torch.testing.assert_close(out[0], loss_r, atol=1e-4, rtol=1e-4)
    if padding is not None:
        valid = ((labels != padding) & (labels - (0 if zero_based else 1) >= 0) & (labels - (0 if zero_based else 1) < C)).sum()
        assert float(out[1]) == float(valid)
    scale = float(dx_r.abs().max())
    # the MMA kernel rounds dlogits to bf16 before the two backward GEMMs (like the layer-by-layer path): 2^-8 relative per term
    tol = 1.0 if impl == 1 else 3.0
    assert float((dx.float() - dx_r).abs().max()) <= 0.01 * tol * scale + 1e-7          # bf16 output
    wscale = float(dw_r.abs().max())
    assert float((dw - dw_r).abs().max()) <= 2e-3 * tol * wscale + 1e-6
    assert float((db - db_r).abs().max()) <= 2e-3 * tol * float(db_r.abs().max()) + 1e-6
    # accumulation: a second launch adds to dw / db
    head_softmax_nll(x16, w16, bias, labels, dw, db, zero_based, padding, True, relu_cols, impl=impl)
    assert float((dw - 2 * dw_r).abs().max()) <= 4e-3 * tol * wscale + 2e-6


def _train(fused, graph, steps=12):
    from zoo.common.nncontext import init_nncontext
    from zoo.feature.common import MiniBatch
    from zoo.models.recommendation import NeuralCF
    from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
    from zoo.pipeline.api.keras.optimizers import Adam
    from zoo.pipeline.engine import TrainingEngine

    os.environ["ZOO_FUSED_HEAD"] = "1" if fused else "0"
    try:
        torch.manual_seed(0)
        ctx = init_nncontext()
        model = NeuralCF(5000, 3000, 5, user_embed=64, item_embed=64, hidden_layers=[128, 64, 64], mf_embed=64)
        eng = TrainingEngine(model, Adam(lr=2e-3), None, ctx)
        eng.use_cuda_graph = graph
        eng.prepare()
        crit = SparseCategoricalCrossEntropy(zero_based_label=False)
        rng = np.random.RandomState(1)
        losses = []
        for _ in range(steps):
            x = torch.from_numpy(np.stack([rng.randint(1, 5001, 2048), rng.randint(1, 3001, 2048)], 1)).cuda()
            y = torch.from_numpy(((x[:, 0].cpu().numpy() * 7 + x[:, 1].cpu().numpy() * 3) % 5 + 1)).cuda()
            losses.append(float(eng.train_step(crit, MiniBatch(x, y, 2048))))
        w = model.linear_last.weight.detach().float().cpu().clone()
        launches = None if eng._graph is None else eng._graph.get("launches")
        return losses, w, launches
    finally:
        os.environ.pop("ZOO_FUSED_HEAD", None)


@pytest.mark.parametrize("graph", [False, True])
def test_engine_step_with_fused_head_matches_layer_by_layer(graph):
    lf, wf, nf = _train(True, graph)
    lu, wu, nu = _train(False, graph)
    np.testing.assert_allclose(lf, lu, rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(wf, wu, atol=3e-3, rtol=5e-2)
    if graph and nf and nu:
        assert nf < nu, (nf, nu)          # the fused head removed launches from the captured step
