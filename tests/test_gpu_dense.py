"""tcgen05 Dense GEMM family vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.to(torch.bfloat16)


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 256, 256), (1000, 128, 256), (4096, 256, 320), (333, 40, 72),
                                   (65536, 256, 256), (512, 5, 128), (130, 16, 8), (2048, 512, 128)])
@pytest.mark.parametrize("act", [None, "relu"])
def test_dense_fwd(M, N, K, act):
    from zoo.ops.dense import dense_forward

    torch.manual_seed(0)
    x = _bf(torch.randn(M, K, device="cuda"))
    w = _bf(torch.randn(N, K, device="cuda") / K ** 0.5)
    b = torch.randn(N, device="cuda")
    y = dense_forward(x, w, b, act)
    ref = x.float() @ w.float().t() + b
    if act == "relu":
        ref = torch.relu(ref)
    torch.cuda.synchronize()
    assert y.shape == (M, N)
    torch.testing.assert_close(y.float(), ref, atol=3e-2, rtol=2e-2)


@pytest.mark.parametrize("act", ["gelu", "sigmoid", "tanh"])
def test_dense_fwd_other_acts(act):
    from zoo.ops.dense import dense_forward, _act_torch

    torch.manual_seed(1)
    x = _bf(torch.randn(300, 96, device="cuda"))
    w = _bf(torch.randn(48, 96, device="cuda") / 10)
    y = dense_forward(x, w, None, act)
    ref = _act_torch(x.float() @ w.float().t(), act)
    torch.testing.assert_close(y.float(), ref, atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 256, 256), (1000, 128, 256), (4096, 64, 128), (333, 40, 72),
                                   (512, 5, 128), (65536, 128, 256)])
@pytest.mark.parametrize("masked", [False, True])
def test_dense_dgrad(M, N, K, masked):
    from zoo.ops.dense import dense_dgrad

    torch.manual_seed(2)
    ldn = (N + 7) // 8 * 8
    dy = torch.zeros(M, ldn, device="cuda", dtype=torch.bfloat16)
    dy[:, :N] = _bf(torch.randn(M, N, device="cuda"))
    dy = dy[:, :N]
    w = _bf(torch.randn(N, K, device="cuda") / N ** 0.5)
    mask = _bf(torch.randn(M, K, device="cuda")) if masked else None
    dx = dense_dgrad(dy, w, mask)
    ref = dy.float() @ w.float()
    if masked:
        ref = ref * (mask.float() > 0)
    torch.testing.assert_close(dx.float(), ref, atol=3e-2, rtol=2e-2)


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 256, 256), (1000, 128, 256), (4096, 64, 128), (333, 40, 72),
                                   (512, 5, 128), (65536, 256, 256), (70000, 16, 264)])
def test_dense_wgrad(M, N, K):
    from zoo.ops.dense import dense_wgrad

    torch.manual_seed(3)
    ldn = (N + 7) // 8 * 8
    dy = torch.zeros(M, ldn, device="cuda", dtype=torch.bfloat16)
    dy[:, :N] = _bf(torch.randn(M, N, device="cuda"))
    dy = dy[:, :N]
    x = _bf(torch.randn(M, K, device="cuda"))
    dw = torch.ones(N, K, device="cuda")  # accumulates on top of existing contents
    dense_wgrad(dy, x, dw)
    ref = dy.float().t() @ x.float() + 1.0
    scale = max(1.0, M ** 0.5)
    torch.testing.assert_close(dw / scale, ref / scale, atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("M,N,K", [(1000, 128, 256), (65536, 256, 256), (4096, 64, 128), (333, 40, 72), (512, 5, 128), (3000, 300, 264)])
def test_dense_wgrad_with_bias_grad(M, N, K):
    from zoo.ops.dense import dense_wgrad

    torch.manual_seed(4)
    ldn = (N + 7) // 8 * 8
    dy = torch.zeros(M, ldn, device="cuda", dtype=torch.bfloat16)
    dy[:, :N] = _bf(torch.randn(M, N, device="cuda"))
    dy = dy[:, :N]
    x = _bf(torch.randn(M, K, device="cuda"))
    dw = torch.ones(N, K, device="cuda")
    db = torch.full((N,), 2.0, device="cuda")
    dense_wgrad(dy, x, dw, db)
    scale = max(1.0, M ** 0.5)
    torch.testing.assert_close(dw / scale, (dy.float().t() @ x.float() + 1.0) / scale, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(db / scale, (dy.float().sum(0) + 2.0) / scale, atol=2e-2, rtol=2e-2)


def test_dense_dgrad_partial_mask():
    from zoo.ops.dense import dense_dgrad

    torch.manual_seed(7)
    M, N, K = 1000, 8, 128
    dy = _bf(torch.randn(M, N, device="cuda"))
    w = _bf(torch.randn(N, K, device="cuda"))
    mask = _bf(torch.randn(M, K, device="cuda"))
    dx = dense_dgrad(dy, w, mask, mask_cols=64)
    ref = dy.float() @ w.float()
    ref[:, :64] = ref[:, :64] * (mask[:, :64].float() > 0)
    torch.testing.assert_close(dx.float(), ref, atol=3e-2, rtol=2e-2)


def _unpack_bits(bits, n):
    sh = torch.arange(32, device=bits.device, dtype=torch.int64)
    b = ((bits.to(torch.int64).unsqueeze(-1) >> sh) & 1).reshape(bits.shape[0], -1)
    return b[:, :n].bool()


@pytest.mark.parametrize("M,N,K", [(1000, 256, 256), (65536, 128, 256), (4096, 64, 128), (333, 96, 72), (2048, 512, 64)])
def test_dense_relu_bits_roundtrip(M, N, K):
    """Forward writes the sign bits of its output; the next layer's dgrad takes them as its ReLU mask."""
    from zoo.ops.dense import dense_dgrad, dense_forward

    torch.manual_seed(11)
    x = _bf(torch.randn(M, K, device="cuda"))
    w = _bf(torch.randn(N, K, device="cuda") / K ** 0.5)
    b = torch.randn(N, device="cuda")
    bits = torch.zeros(M, (N + 31) // 32, dtype=torch.int32, device="cuda")
    y = dense_forward(x, w, b, "relu", relu_bits=bits)
    torch.cuda.synchronize()
    assert torch.equal(_unpack_bits(bits, N), y > 0)
    # consumer layer: N -> 40 features; its dgrad produces dY of width N masked by the bits
    w2 = _bf(torch.randn(40, N, device="cuda") / N ** 0.5)
    dz = _bf(torch.randn(M, 40, device="cuda"))
    dy = dense_dgrad(dz, w2, mask_bits=bits)
    ref = (dz.float() @ w2.float()) * (y > 0)
    torch.testing.assert_close(dy.float(), ref, atol=3e-2, rtol=2e-2)
    if N >= 64:
        dy2 = dense_dgrad(dz, w2, mask_bits=bits, mask_cols=32)
        ref2 = dz.float() @ w2.float()
        ref2[:, :32] = ref2[:, :32] * (y[:, :32] > 0)
        torch.testing.assert_close(dy2.float(), ref2, atol=3e-2, rtol=2e-2)


def test_dense_fwd_into_column_window():
    """`out` may be a column window of a wider buffer (fused concat): neighbouring columns stay untouched."""
    from zoo.ops.dense import dense_forward

    torch.manual_seed(12)
    M, N, K = 777, 72, 64
    x = _bf(torch.randn(M, K, device="cuda"))
    w = _bf(torch.randn(N, K, device="cuda") / 8)
    buf = torch.full((M, 160), 7.0, device="cuda", dtype=torch.bfloat16)
    dense_forward(x, w, None, None, out=buf[:, 8:8 + N])
    torch.testing.assert_close(buf[:, 8:8 + N].float(), x.float() @ w.float().t(), atol=3e-2, rtol=2e-2)
    assert bool((buf[:, :8] == 7).all()) and bool((buf[:, 8 + N:] == 7).all())


def test_colsum():
    from zoo.ops.dense import colsum

    for rows, cols in [(5000, 72), (65536, 256), (777, 8), (4096, 64), (3000, 520), (100, 5)]:
        ld = (cols + 7) // 8 * 8
        dy = _bf(torch.randn(rows, ld, device="cuda"))[:, :cols]
        out = torch.zeros(cols, device="cuda")
        colsum(dy, out)
        torch.testing.assert_close(out, dy.float().sum(0), atol=2e-2, rtol=2e-3)


def test_fused_mlp_autograd_matches_torch():
    from zoo.ops.dense import FusedMLP

    torch.manual_seed(4)
    mlp = FusedMLP(64, [128, 64, 32]).cuda()
    x = torch.randn(512, 64, device="cuda")
    xk = x.to(torch.bfloat16).requires_grad_(True)
    y = mlp(xk)
    y.float().pow(2).mean().backward()
    # reference in fp32 with the same (bf16-rounded) weights
    xr = x.to(torch.bfloat16).float().requires_grad_(True)
    h = xr
    for layer in mlp:
        h = torch.relu(h @ layer.weight.detach().to(torch.bfloat16).float().t() + layer.bias.detach())
    h.pow(2).mean().backward()
    torch.testing.assert_close(y.float(), h, atol=3e-2, rtol=3e-2)
    torch.testing.assert_close(xk.grad.float(), xr.grad, atol=2e-3, rtol=5e-2)
    assert mlp[0].weight.grad is not None and torch.isfinite(mlp[0].weight.grad).all()


@pytest.mark.parametrize("M,dims,window", [(1000, (256, 256, 128, 64), False), (130, (64, 64, 64, 64), False), (4099, (128, 192, 128, 64), True),
                                           (65536, (256, 256, 128, 64), True)])
def test_fused_tower_matches_layer_by_layer_and_fp32(M, dims, window):
    """K3 whole-tower kernel (csrc/gemm/mlp3_tcgen05.cu): forward values, the activations saved for backward and every gradient
    against the layer-by-layer kernels (same arithmetic: bit-equal forward expected up to accumulation order) and an fp32 torch tower."""
    import os

    from zoo.ops.dense import FusedMLP

    torch.manual_seed(M + sum(dims))
    dev = "cuda"
    mlp = FusedMLP(dims[0], list(dims[1:]), activation="relu").to(dev)
    with torch.no_grad():
        for l in mlp:
            l.bias.uniform_(-0.1, 0.1)
    x = (torch.randn(M, dims[0], device=dev) * 0.5).to(torch.bfloat16)
    wide = torch.zeros(M, dims[3] + 64, dtype=torch.bfloat16, device=dev) if window else None

    def run(fused):
        os.environ["ZOO_FUSED_TOWER"] = "1" if fused else "0"
        try:
            for p in mlp.parameters():
                p.grad = None
            xi = x.clone().requires_grad_(True)
            out = None if wide is None else wide[:, :dims[3]]
            y = mlp(xi, out=out) if out is not None else mlp(xi)
            (y.float() * torch.linspace(-1, 1, dims[3], device=dev)).sum().backward()
            return y.detach().float().clone(), xi.grad.float().clone(), [p.grad.float().clone() for p in mlp.parameters()]
        finally:
            os.environ.pop("ZOO_FUSED_TOWER", None)

    yf, dxf, gf = run(True)
    yu, dxu, gu = run(False)
    torch.testing.assert_close(yf, yu, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(dxf, dxu, atol=3e-2, rtol=5e-2)
    for a, b in zip(gf, gu):
        torch.testing.assert_close(a, b, atol=2e-2 * max(1.0, float(b.abs().max())), rtol=5e-2)
    # fp32 reference of the same tower (bf16 inputs / weights, fp32 math)
    h = x.float()
    for l in mlp:
        h = torch.relu(h @ l.weight.detach().to(torch.bfloat16).float().t() + l.bias.detach().float())
    assert float((yf - h).abs().max()) <= 0.03 * max(1.0, float(h.abs().max()))
    if window:
        assert float(wide[:, dims[3]:].abs().max()) == 0.0          # the kernel wrote only its column window
