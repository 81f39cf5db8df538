"""Embedding / optimizer / loss kernels vs plain PyTorch fp32 references."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("D", [128, 20, 64])
@pytest.mark.parametrize("shadow", [False, True])
def test_multi_lookup_forward(D, shadow):
    from zoo.ops.embedding import EmbeddingTable, LookupSpec, multi_lookup

    torch.manual_seed(0)
    ta = EmbeddingTable(1000, D, shadow_bf16=shadow).cuda()
    tb = EmbeddingTable(500, D, shadow_bf16=shadow).cuda()
    ids = torch.stack([torch.randint(0, 1000, (4096,)), torch.randint(0, 500, (4096,))], 1).cuda()
    out = multi_lookup(ids, [LookupSpec(ta, 0), LookupSpec(tb, 1)])
    ref = torch.cat([ta.weight[ids[:, 0]], tb.weight[ids[:, 1]]], 1)
    torch.testing.assert_close(out.float(), ref.to(torch.bfloat16).float(), atol=1e-2, rtol=1e-2)


def test_lookup_mul_and_bag():
    from zoo.ops.embedding import EmbeddingTable, LookupSpec, lookup_mul, multi_lookup

    torch.manual_seed(1)
    ta = EmbeddingTable(300, 64).cuda()
    tb = EmbeddingTable(200, 64).cuda()
    ids = torch.stack([torch.randint(0, 300, (1024,)), torch.randint(0, 200, (1024,))], 1).cuda()
    out = lookup_mul(ids, ta, 0, tb, 1)
    ref = ta.weight[ids[:, 0]] * tb.weight[ids[:, 1]]
    torch.testing.assert_close(out.float(), ref, atol=1e-3, rtol=2e-2)
    bag_ids = torch.randint(-1, 300, (512, 6)).cuda()
    outb = multi_lookup(bag_ids, [LookupSpec(ta, 0, bag=6)])
    w = torch.cat([ta.weight, torch.zeros(1, 64, device="cuda")])
    refb = w[bag_ids.clamp(min=-1)].sum(1)
    torch.testing.assert_close(outb.float(), refb, atol=2e-3, rtol=2e-2)


@pytest.mark.parametrize("opt", ["sgd", "adam", "adagrad"])
@pytest.mark.parametrize("rows,B", [(50, 4096), (100000, 4096)])
def test_sparse_update_matches_dense_reference(opt, rows, B):
    """Sort-free dedup + fused sparse optimizer == torch optimizer fed the dense (summed) gradient,
    restricted to touched rows (lazy semantics)."""
    from zoo.ops.embedding import EmbeddingTable, lookup

    torch.manual_seed(2)
    D = 32
    t = EmbeddingTable(rows, D, optimizer=opt, lr=0.05).cuda()
    w0 = t.weight.detach().clone()
    ids = torch.randint(0, rows, (B, 1)).cuda()
    target = torch.randn(B, D, device="cuda")
    ref_w = w0.clone().requires_grad_(True)
    ref_opt = {"sgd": lambda p: torch.optim.SGD(p, lr=0.05), "adam": lambda p: torch.optim.Adam(p, lr=0.05),
               "adagrad": lambda p: torch.optim.Adagrad(p, lr=0.05, eps=1e-8)}[opt]([ref_w])
    for step in range(3):
        out = lookup(ids, t)
        (out.float() * target).sum().backward()
        ref_opt.zero_grad()
        g = torch.zeros_like(ref_w)
        g.index_add_(0, ids[:, 0], target.to(torch.bfloat16).float())
        touched = torch.zeros(rows, dtype=torch.bool, device="cuda")
        touched[ids[:, 0]] = True
        before = ref_w.detach().clone()
        ref_w.grad = g
        ref_opt.step()
        with torch.no_grad():  # lazy: untouched rows keep their value
            ref_w[~touched] = before[~touched]
    torch.cuda.synchronize()
    # every step touches the same rows, so lazy == dense on them
    torch.testing.assert_close(t.weight.detach(), ref_w.detach(), atol=2e-3, rtol=2e-2)
    assert int((t.rep_of_row != 2**31 - 1).sum()) == 0, "scratch map must be restored"
    if t.accum is not None:            # per-table path
        assert float(t.accum.abs().sum()) == 0.0 and int(t.dup_flag.sum()) == 0
    else:                              # grouped path keeps its scratch in the lookup plan
        import zoo.ops.embedding_group as G

        assert G.scratch_is_clean()


def test_dense_grad_mode():
    from zoo.ops.embedding import EmbeddingTable, lookup

    torch.manual_seed(3)
    t = EmbeddingTable(64, 20, sparse=False).cuda()
    ids = torch.randint(0, 64, (2000, 1)).cuda()
    tgt = torch.randn(2000, 20, device="cuda")
    (lookup(ids, t).float() * tgt).sum().backward()
    ref = torch.zeros(64, 20, device="cuda").index_add_(0, ids[:, 0], tgt.to(torch.bfloat16).float())
    torch.testing.assert_close(t.weight.grad, ref, atol=1e-2, rtol=1e-2)


@pytest.mark.parametrize("name", ["adam", "adamw", "sgd", "adagrad", "rmsprop"])
def test_fused_optimizer_matches_torch(name):
    from zoo.parallel.flat import FusedOptimizer, OptConfig

    torch.manual_seed(4)
    net = torch.nn.Sequential(torch.nn.Linear(33, 17), torch.nn.ReLU(), torch.nn.Linear(17, 3)).cuda()
    ref = torch.nn.Sequential(torch.nn.Linear(33, 17), torch.nn.ReLU(), torch.nn.Linear(17, 3)).cuda()
    ref.load_state_dict(net.state_dict())
    mk = {"adam": lambda p: torch.optim.Adam(p, lr=1e-2, weight_decay=1e-3),
          "adamw": lambda p: torch.optim.AdamW(p, lr=1e-2, weight_decay=1e-2),
          "sgd": lambda p: torch.optim.SGD(p, lr=1e-2, momentum=0.9, dampening=0.0),
          "adagrad": lambda p: torch.optim.Adagrad(p, lr=1e-2),
          "rmsprop": lambda p: torch.optim.RMSprop(p, lr=1e-2)}[name]
    ropt = mk(ref.parameters())
    fopt = FusedOptimizer.build(net, OptConfig.from_torch(mk(net.parameters())))
    for _ in range(4):
        x = torch.randn(64, 33, device="cuda")
        for m_, o in ((net, None), (ref, ropt)):
            if o is not None:
                o.zero_grad()
            m_(x).pow(2).mean().backward()
        ropt.step()
        fopt.step()
    for p, q in zip(net.parameters(), ref.parameters()):
        torch.testing.assert_close(p, q, atol=2e-5, rtol=2e-4)
    assert float(fopt.arena.flat_g.abs().sum()) == 0.0
    torch.testing.assert_close(fopt.arena.flat_p16.float(), fopt.arena.flat_p, atol=1e-2, rtol=1e-2)


def test_l2_clip_and_sumsq():
    from zoo.parallel.flat import FusedOptimizer, OptConfig

    net = torch.nn.Linear(1000, 100).cuda()
    fopt = FusedOptimizer.build(net, OptConfig("sgd", lr=1.0, b1=0.0))
    fopt.set_gradient_clipping_by_l2_norm(0.5)
    fopt.arena.flat_g.normal_()
    g0 = fopt.arena.flat_g.clone()
    p0 = fopt.arena.flat_p.clone()
    fopt.step()
    upd = p0 - fopt.arena.flat_p
    torch.testing.assert_close(fopt.sumsq[0], (g0 ** 2).sum(), rtol=1e-4, atol=1e-2)
    torch.testing.assert_close(upd.norm(), torch.tensor(0.5, device="cuda"), atol=1e-3, rtol=1e-3)


@pytest.mark.parametrize("C", [2, 5, 100, 1000])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_softmax_nll(C, dtype):
    from zoo.ops.loss import softmax_nll, softmax_nll_reference

    torch.manual_seed(5)
    B = 777
    logits = (torch.randn(B, C, device="cuda") * 3).to(dtype).requires_grad_(True)
    labels = torch.randint(1, C + 1, (B,), device="cuda")
    loss = softmax_nll(logits, labels, zero_based_label=False)
    loss.backward()
    lr_ = logits.detach().float().requires_grad_(True)
    ref, _ = softmax_nll_reference(lr_, labels, zero_based_label=False)
    ref.backward()
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    torch.testing.assert_close(loss, ref, atol=tol, rtol=tol)
    torch.testing.assert_close(logits.grad.float(), lr_.grad, atol=tol / 10, rtol=5e-2)


def test_softmax_nll_padding_and_weights():
    from zoo.ops.loss import softmax_nll, softmax_nll_reference

    torch.manual_seed(6)
    logits = torch.randn(100, 7, device="cuda", requires_grad=True)
    labels = torch.randint(0, 7, (100,), device="cuda")
    labels[::5] = -1
    w = torch.rand(7, device="cuda") + 0.5
    loss = softmax_nll(logits, labels, padding_value=-1, class_weights=w)
    loss.backward()
    lr_ = logits.detach().clone().requires_grad_(True)
    ref, _ = softmax_nll_reference(lr_, labels, padding_value=-1, class_weights=w)
    ref.backward()
    torch.testing.assert_close(loss, ref, atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(logits.grad, lr_.grad, atol=1e-6, rtol=1e-3)


def test_sigmoid_bce():
    from zoo.ops.loss import sigmoid_bce, sigmoid_bce_reference

    x = torch.randn(5000, device="cuda", requires_grad=True)
    y = (torch.rand(5000, device="cuda") > 0.5).float()
    l = sigmoid_bce(x, y)
    l.backward()
    xr = x.detach().clone().requires_grad_(True)
    r = sigmoid_bce_reference(xr, y)
    r.backward()
    torch.testing.assert_close(l, r, atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(x.grad, xr.grad, atol=1e-7, rtol=1e-3)


def test_mean_pooled_padded_bags_match_cpu():
    """Mean pooling divides by the number of PRESENT ids (padding = -1) in forward AND backward, on the GPU kernels and on the
    plain-torch CPU path alike."""
    import copy

    from zoo.ops.embedding import EmbeddingTable, lookup

    torch.manual_seed(0)
    rows, D, B, bag = 50, 16, 64, 4
    t_cpu = EmbeddingTable(rows, D, sparse=False)
    t_gpu = copy.deepcopy(t_cpu).cuda()
    ids = torch.randint(0, rows, (B, bag))
    ids[torch.rand(B, bag) < 0.4] = -1
    ids[:, 0] = torch.randint(0, rows, (B,))               # at least one present id per bag
    w = torch.randn(B, D)
    out_c = lookup(ids, t_cpu, 0, bag, mean=True)
    (out_c * w).sum().backward()
    out_g = lookup(ids.cuda(), t_gpu, 0, bag, mean=True)
    (out_g.float() * w.cuda()).sum().backward()
    torch.testing.assert_close(out_g.float().cpu(), out_c, atol=2e-2, rtol=2e-2)
    g_gpu = t_gpu.weight.grad if t_gpu.weight.grad is not None else t_gpu.weight._zoo_grad
    torch.testing.assert_close(g_gpu.float().cpu(), t_cpu.weight.grad, atol=3e-2, rtol=3e-2)
