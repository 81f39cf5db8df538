"""FP8 (E4M3) GEMM on tcgen05 (kind::f8f6f4) against the CPU restatement of the same arithmetic (dequantised operands, fp32
accumulation), and end to end against the fp32 model."""
import copy

import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,K,N,act", [(256, 128, 64, None), (1000, 200, 96, "relu"), (4096, 256, 256, "relu"), (77, 40, 5, "sigmoid")])
@pytest.mark.parametrize("chain", [False, True])
def test_fp8_linear_matches_cpu_restatement(M, K, N, act, chain):
    from zoo.ops.quant import FP8Linear, FP8Tensor

    torch.manual_seed(0)
    w, b = torch.randn(N, K) * 0.2, torch.randn(N) * 0.1
    x = torch.randn(M, K)
    sx = float(x.abs().max()) / 448.0
    lin = FP8Linear(w, b, act, in_scale=sx, next_scale=(0.05 if chain else None))
    want = lin(x)                                     # CPU path: the oracle
    got = copy.deepcopy(lin).cuda()(x.cuda())
    if chain:
        assert isinstance(got, FP8Tensor) and got.q.shape == want.q.shape
        # both sides round the same fp32 values to E4M3; accumulation order may move a value across a rounding boundary (1 ulp = 12.5 %)
        diff = (got.float().cpu() - want.float()).abs()
        assert float((diff > 0.13 * want.float().abs() + 1e-3).float().mean()) < 0.01
        assert float(diff.mean()) < 2e-3
    else:
        torch.testing.assert_close(got.float().cpu(), want, atol=2e-2, rtol=2e-2)


def test_quantized_ncf_tower_close_to_fp32():
    from zoo.models.recommendation import NeuralCF
    from zoo.ops.quant import quantize_linear_layers

    torch.manual_seed(0)
    m = NeuralCF(500, 300, 5, user_embed=64, item_embed=64, hidden_layers=[128, 64, 32], mf_embed=32, sparse=False).cuda().eval()
    x = torch.stack([torch.randint(1, 501, (2048,)), torch.randint(1, 301, (2048,))], 1).cuda()
    with torch.no_grad():
        want = m(x).float()
        q = quantize_linear_layers(copy.deepcopy(m), x)
        got = q(x).float()
    assert sum(type(k).__name__ == "FP8Linear" for k in q.modules()) == 4
    assert float((got - want).abs().max()) < 0.03
    assert float((got.argmax(1) == want.argmax(1)).float().mean()) > 0.9
