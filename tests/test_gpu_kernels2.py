"""K7 (TCN conv), K13 (serving preprocess), K14 (top-k), K15 (recommend top-N) against CPU oracles."""
import copy

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,Cin,Cout,T,K,d", [(3, 5, 7, 50, 3, 1), (4, 16, 16, 130, 7, 4), (2, 30, 30, 257, 3, 64), (2, 8, 20, 40, 2, 8),
                                              (2, 64, 64, 256, 7, 16), (2, 96, 80, 100, 3, 2), (3, 32, 48, 64, 5, 1), (2, 16, 24, 300, 2, 200)])
@pytest.mark.parametrize("with_res", [False, True])
@pytest.mark.parametrize("mma", [True, False])
def test_causal_conv_fwd_bwd(B, Cin, Cout, T, K, d, with_res, mma, monkeypatch):
    """Both forms against the fp32 torch restatement: the TF32 tensor-core kernels (csrc/ops/tcn_mma.cu; tolerance of TF32 operands,
    the precision cuDNN convolutions use by default) where they apply, and the fp32 CUDA-core kernels (ZOO_TCN_MMA=0)."""
    from zoo.ops.tcn import causal_conv1d, causal_conv1d_reference, mma_path

    monkeypatch.setenv("ZOO_TCN_MMA", "1" if mma else "0")
    tf32 = mma_path(Cin, Cout, T, K, d)
    if mma and not tf32:
        pytest.skip("shape runs on the CUDA-core kernels either way")
    assert tf32 == mma
    torch.manual_seed(0)
    x = torch.randn(B, Cin, T)
    w = torch.randn(Cout, Cin, K) * 0.3
    b = torch.randn(Cout) * 0.1
    res = torch.randn(B, Cout, T) if with_res else None
    leaves = [t.clone().requires_grad_(True) for t in (x, w, b)] + ([res.clone().requires_grad_(True)] if with_res else [None])
    ref = causal_conv1d_reference(leaves[0], leaves[1], leaves[2], d, True, leaves[3])
    wgt = torch.linspace(-1, 1, ref.numel()).reshape(ref.shape)
    (ref * wgt).sum().backward()
    dl = [t.detach().cuda().requires_grad_(True) for t in (x, w, b)] + ([res.cuda().requires_grad_(True)] if with_res else [None])
    out = causal_conv1d(dl[0], dl[1], dl[2], d, True, dl[3])
    (out * wgt.cuda()).sum().backward()
    if tf32:
        scale = float(ref.detach().abs().max())
        torch.testing.assert_close(out.detach().cpu(), ref.detach(), atol=4e-3 * scale, rtol=1e-2)
        # Gradients: a ReLU whose pre-activation sits within TF32 rounding of zero may fall on the other side of the kink, and one
        # flipped element moves dW by several per cent of its largest entry.  So the oracle takes the masks from the KERNEL's forward
        # output (as the kernel's backward does) and restates the three gradient contractions in fp32.
        o = out.detach().cpu()
        if with_res:
            dmid = wgt * (o > 0)
            g = dmid * ((o - res) > 0)
        else:
            dmid, g = None, wgt * (o > 0)
        xp = torch.nn.functional.pad(x, ((K - 1) * d, 0))
        want = [torch.nn.functional.conv_transpose1d(g, w, dilation=d)[:, :, (K - 1) * d:],
                torch.stack([torch.einsum("bot,bit->oi", g, xp[:, :, j * d:j * d + T]) for j in range(K)], -1), g.sum((0, 2)), dmid]
        for a, r in zip(dl, want):
            if a is not None:
                err = (a.grad.cpu() - r).abs().max() / r.abs().max().clamp_min(1e-6)
                assert err < 5e-3, float(err)
    else:
        torch.testing.assert_close(out.detach().cpu(), ref.detach(), atol=1e-4, rtol=1e-4)
        for a, r in zip(dl, leaves):
            if a is not None:
                torch.testing.assert_close(a.grad.cpu(), r.grad, atol=2e-3, rtol=2e-3)


def test_causal_conv_tf32_dropout_matches_cuda_core_mask(monkeypatch):
    """The two forms draw the same dropout mask from the same seed (one hash of the output index)."""
    from zoo.ops.tcn import causal_conv1d

    torch.manual_seed(1)
    x, w = torch.randn(2, 32, 96).cuda(), (torch.randn(32, 32, 3) * 0.2).cuda()
    monkeypatch.setenv("ZOO_TCN_MMA", "1")
    a = causal_conv1d(x, w, None, 2, True, None, drop_p=0.3, training=True, seed=11)
    monkeypatch.setenv("ZOO_TCN_MMA", "0")
    b = causal_conv1d(x, w, None, 2, True, None, drop_p=0.3, training=True, seed=11)
    both = (a != 0) & (b != 0)
    assert ((a != 0) != (b != 0)).float().mean() < 1e-3          # only ReLU kinks may differ
    torch.testing.assert_close(a[both], b[both], atol=1e-2, rtol=1e-2)


def test_tcn_block_dropout_and_training():
    from zoo.ops.tcn import TemporalConvNet, causal_conv1d

    torch.manual_seed(0)
    x = torch.randn(4, 6, 64).cuda()
    w = torch.randn(9, 6, 3).cuda()
    y = causal_conv1d(x, w, None, 2, True, None, drop_p=0.5, training=True, seed=7)
    y0 = causal_conv1d(x, w, None, 2, True, None)
    kept = y != 0
    frac = 1 - kept.float().sum() / (y0 != 0).float().sum()
    assert 0.35 < frac.item() < 0.65
    torch.testing.assert_close(y[kept], 2 * y0[kept])
    y2 = causal_conv1d(x, w, None, 2, True, None, drop_p=0.5, training=True, seed=7)
    assert torch.equal(y, y2)                       # same seed, same mask
    net = TemporalConvNet(6, [8, 8, 4], kernel_size=3, dropout=0.1)
    ref = copy.deepcopy(net).eval()
    net.cuda().eval()
    torch.testing.assert_close(net(x).cpu(), ref(x.cpu()), atol=1e-4, rtol=1e-3)
    net.train()
    opt = torch.optim.Adam(net.parameters(), 1e-2)
    tgt = torch.randn(4, 4, 64).cuda()
    l0 = None
    for _ in range(40):
        opt.zero_grad()
        loss = (net(x) - tgt).pow(2).mean()
        loss.backward()
        opt.step()
        l0 = l0 or loss.item()
    assert loss.item() < l0


@pytest.mark.parametrize("chw,swap,dtype", [(True, False, torch.float32), (False, True, torch.float32), (True, True, torch.bfloat16)])
def test_serving_preproc(chw, swap, dtype):
    from zoo.ops.serving_ops import preprocess_images, preprocess_reference

    rng = np.random.RandomState(0)
    imgs = [rng.randint(0, 256, (h, w, 3), dtype=np.uint8) for h, w in [(300, 400), (224, 224), (97, 513), (600, 350), (256, 256)]]
    mean, std = [123.0, 117.0, 104.0], [58.0, 57.0, 57.5]
    out = preprocess_images(imgs, (224, 224), (256, 256), mean, std, swap, chw, dtype, device="cuda")
    ref = preprocess_reference(imgs, (224, 224), (256, 256), mean, std, swap, chw)
    tol = dict(atol=3e-2, rtol=2e-2) if dtype == torch.bfloat16 else dict(atol=2e-4, rtol=1e-4)
    torch.testing.assert_close(out.float().cpu(), ref, **tol)
    # geometry agrees with OpenCV's INTER_LINEAR (cv2 uses fixed-point weights for u8: <= 1 grey level apart)
    import cv2

    o2 = preprocess_images(imgs[:1], (200, 320), None, None, None, False, False, torch.float32, device="cuda")[0].cpu().numpy()
    c2 = cv2.resize(imgs[0], (320, 200), interpolation=cv2.INTER_LINEAR).astype(np.float32)
    assert np.abs(o2 - c2).max() <= 1.01


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n,k,softmax", [(1000, 5, True), (1000, 1, False), (37, 8, True), (5, 3, False)])
def test_serving_topk(dtype, n, k, softmax):
    from zoo.ops.serving_ops import topk

    torch.manual_seed(0)
    x = (torch.randn(67, n) * 3).to(dtype).cuda()
    idx, val = topk(x, k, softmax)
    xf = x.float().cpu()
    ref = (torch.softmax(xf, -1) if softmax else xf).topk(k, dim=-1)
    torch.testing.assert_close(val.cpu(), ref.values, atol=1e-5, rtol=1e-4)
    # indices may differ only between exactly tied values
    got = (torch.softmax(xf, -1) if softmax else xf).gather(1, idx.cpu().long())
    torch.testing.assert_close(got, ref.values, atol=1e-5, rtol=1e-4)


def test_recommend_topn():
    from zoo.ops.recommend import pair_score, segmented_topn_order

    torch.manual_seed(0)
    P, C, U, N = 20000, 5, 300, 4
    probs = torch.softmax(torch.randn(P, C), -1)
    users = torch.randint(0, U, (P,))
    cls_, p, key = pair_score(probs.cuda())
    rc, rp, rkey = pair_score(probs)
    assert torch.equal(cls_.cpu(), rc) and torch.equal(p.cpu(), rp) and torch.equal(key.cpu(), rkey)
    sel = segmented_topn_order(users.cuda(), key, N).cpu()
    ref = segmented_topn_order(users, rkey, N)
    assert torch.equal(sel, ref)
    # semantic check against a python groupby
    by = {}
    for i in range(P):
        by.setdefault(int(users[i]), []).append((-int(rc[i]), -float(rp[i]), i))
    want = [i for u in sorted(by) for _, _, i in sorted(by[u])[:N]]
    assert sel.tolist() == want


def test_nvjpeg_batch_decode_and_preprocess_match_cv2():
    """JPEG bytes -> nvJPEG batched decode on the GPU -> K13 must match cv2.imdecode + the CPU preprocessing oracle (decoders may
    differ by a couple of grey levels in their IDCT / chroma upsampling)."""
    import cv2
    import numpy as np

    from zoo.ops.serving_ops import JpegBatchDecoder, preprocess_jpegs, preprocess_reference

    if not JpegBatchDecoder.available():
        pytest.skip("libnvjpeg not available")
    rng = np.random.RandomState(0)
    imgs = []
    for (h, w) in [(224, 224), (180, 260), (300, 200), (64, 64)]:
        base = cv2.GaussianBlur(cv2.resize(rng.randint(0, 256, (h // 16 + 1, w // 16 + 1, 3), dtype=np.uint8), (w, h), interpolation=cv2.INTER_CUBIC),
                (9, 9), 0)   # smooth content
        imgs.append(base)
    jpgs = [cv2.imencode(".jpg", im, [cv2.IMWRITE_JPEG_QUALITY, 92])[1].tobytes() for im in imgs]
    dec = JpegBatchDecoder("cuda:0")
    mean, std = [123.68, 116.779, 103.939], [58.393, 57.12, 57.375]
    for _ in range(2):                              # twice: the decoder state is reused
        got = preprocess_jpegs(jpgs, (112, 112), dec, None, mean, std, swap_rb=True, chw=True, dtype=torch.float32)
        torch.cuda.synchronize()
    want = preprocess_reference([cv2.imdecode(np.frombuffer(j, np.uint8), cv2.IMREAD_COLOR) for j in jpgs], (112, 112), None, mean, std, True, True)
    diff = (got.cpu() - want).abs() * 58.0        # back to grey levels
    assert float(diff.mean()) < 2.5 and float(diff.max()) < 40, (float(diff.mean()), float(diff.max()))   # chroma upsampling differs between decoders
    dec.close()
