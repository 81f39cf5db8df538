"""Grouped embedding kernels (csrc/ops/embedding_group.cu) and the push all-reduce (csrc/comm/ar_push.cu) against plain fp32
PyTorch restatements.  Multi-rank cases run all 'ranks' on ONE GPU: none of these kernels waits for another launch (the
device barrier between routing / staging and the owner passes is a separate kernel, replaced here by stream order), so the
routing, ownership and reduction logic of world = 2 / 4 / 8 is exercised on a single device."""
import copy
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
INT_MAX = 2**31 - 1


def _zipf(n, rows, alpha=1.05):
    u = torch.rand(n, dtype=torch.float64)
    a = 1.0 - alpha
    x = ((u * ((rows + 1.0) ** a - 1.0)) + 1.0) ** (1.0 / a)
    return x.floor().clamp_(1, rows).long() - 1


def _adam_ref(w, g, touched, lr=1e-2, step=1.0):
    m = 0.1 * g
    v = 0.001 * g * g
    bc1, bc2 = 1 - 0.9 ** step, 1 - 0.999 ** step
    new = w - lr * (m / bc1) / ((v / bc2).sqrt() + 1e-8)
    return torch.where(touched[:, None], new, w)


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("dist_kind", ["uniform", "zipf"])
def test_route_and_grouped_update_emulated(world, dist_kind):
    """Two id columns, two tables per column (one with a product factor): route -> elect / accumulate / apply must equal a dense
    index_add gradient + one Adam step on the touched rows of every table, for every owner."""
    import zoo.ops.embedding_group as G
    from zoo.native import lib as nl

    lib = G._lib()
    torch.manual_seed(3)
    W, B = world, 600
    rows = [1009, 257]                    # column 0 / column 1 tables
    dims = [(32, 16), (32, 16)]           # (table 0, table 1) widths per column
    st = torch.cuda.current_stream().cuda_stream
    if dist_kind == "zipf":
        ids = [torch.stack([_zipf(B, rows[0]), _zipf(B, rows[1])], 1).cuda() for _ in range(W)]
    else:
        ids = [torch.stack([torch.randint(0, rows[0], (B,)), torch.randint(0, rows[1], (B,))], 1).cuda() for _ in range(W)]
    ld = 32 + 32 + 16 + 8                 # one gradient buffer per source: [t00 | t10 | t01 (and t11 share) | pad]
    dout = [torch.randn(B, ld, device="cuda").to(torch.bfloat16) for _ in range(W)]
    fa = [torch.randn(B, 16, device="cuda").to(torch.bfloat16) for _ in range(W)]     # factor rows for column 0's second table
    fb = [torch.randn(B, 16, device="cuda").to(torch.bfloat16) for _ in range(W)]     # ... and column 1's
    col_off = {(0, 0): 0, (1, 0): 32, (0, 1): 64, (1, 1): 64}                           # (column, table) -> window offset
    full = {(c, t): torch.randn(rows[c], dims[c][t], device="cuda") for c in range(2) for t in range(2)}
    scale = 1.0 / W
    # ---- reference ----------------------------------------------------------------------------------------------------------
    ref = {}
    for c in range(2):
        touched = torch.zeros(rows[c], dtype=torch.bool, device="cuda")
        for s in range(W):
            touched[ids[s][:, c]] = True
        for t in range(2):
            D = dims[c][t]
            g = torch.zeros(rows[c], D, device="cuda")
            for s in range(W):
                gr = dout[s][:, col_off[(c, t)]:col_off[(c, t)] + D].float()
                if t == 1:
                    gr = gr * (fb[s] if c == 0 else fa[s]).float()
                g.index_add_(0, ids[s][:, c], gr * scale)
            ref[(c, t)] = _adam_ref(full[(c, t)], g, touched)
    # ---- route (world > 1): every source pushes its (local row, entry) pairs into the owners' mailboxes --------------------------
    cap = B
    pairs = [[torch.zeros(W * cap, 2, dtype=torch.int32, device="cuda") for _ in range(W)] for _ in range(2)]
    counts = [[torch.zeros(8, dtype=torch.int32, device="cuda") for _ in range(W)] for _ in range(2)]
    if W > 1:
        for s in range(W):
            p = G.EgRouteParams()
            p.ids, p.id_stride, p.B, p.world, p.rank, p.cap, p.ngroups = ids[s].data_ptr(), 2, B, W, s, cap, 2
            ctr = [torch.zeros(9, dtype=torch.int32, device="cuda") for _ in range(2)]
            for c in range(2):
                for r in range(W):
                    p.g[c].pairs[r] = pairs[c][r].data_ptr()
                    p.g[c].counts[r] = counts[c][r].data_ptr()
                p.g[c].counters, p.g[c].rows, p.g[c].id_col, p.g[c].bag = ctr[c].data_ptr(), rows[c], c, 1
            nl.check(lib.zoo_eg_route(C.byref(p), st), "zoo_eg_route")
            torch.cuda.synchronize()
            assert all(int(x.sum()) == 0 for x in ctr)                     # local counters restored
        for c in range(2):
            assert sum(int(counts[c][r][:W].sum()) for r in range(W)) == W * B
    # ---- owner passes -------------------------------------------------------------------------------------------------------------
    for r in range(W):
        p = G.EgParams()
        p.ngroups = 2
        keep, reps, dups = [], [], []
        shards = {}
        for c in range(2):
            rl = (rows[c] + W - 1) // W
            e = p.g[c]
            rep = torch.full((rl,), INT_MAX, dtype=torch.int32, device="cuda")
            maxE = B if W == 1 else W * cap
            dup = torch.zeros(2 * maxE, dtype=torch.int32, device="cuda")
            win = torch.full((2 * maxE + 2,), -9, dtype=torch.int32, device="cuda")
            wc = torch.full((2,), 77, dtype=torch.int32, device="cuda")
            e.ids, e.id_stride, e.id_col, e.bag, e.E_local = ids[r].data_ptr(), 2, c, 1, B
            e.pairs, e.counts = pairs[c][r].data_ptr(), counts[c][r].data_ptr()
            e.rep_of_row, e.dup_flag, e.win_list, e.win_count = rep.data_ptr(), dup.data_ptr(), win.data_ptr(), wc.data_ptr()
            e.rows, e.cap, e.world, e.rank, e.ntabs, e.max_entries = rows[c], cap, W, r, 2, maxE
            keep += [win, wc]
            reps.append(rep)
            dups.append(dup)
            for t in range(2):
                D = dims[c][t]
                w = torch.zeros(rl, D, device="cuda")
                src_rows = full[(c, t)][r::W]
                w[: src_rows.shape[0]] = src_rows
                w16 = torch.zeros(rl, D, dtype=torch.bfloat16, device="cuda")
                ea, es = torch.zeros_like(w), torch.zeros_like(w)
                accum = torch.zeros(maxE, D, device="cuda")
                hyper = torch.tensor([1e-2, 0.9, 0.999, 1e-8, 0.0, 0.0] + [0.0] * 10, device="cuda")     # step 0 -> bumped to 1
                tt = e.t[t]
                tt.w, tt.w16, tt.m, tt.v, tt.accum, tt.hyper = w.data_ptr(), w16.data_ptr(), ea.data_ptr(), es.data_ptr(), accum.data_ptr(), hyper.data_ptr()
                tt.D, tt.opt, tt.grad_scale, tt.bump_step, tt.ldg, tt.ldmul = D, 1, scale, 1, ld, 16
                for s in range(W):
                    tt.dout[s] = dout[s].data_ptr() + 2 * col_off[(c, t)]
                    if t == 1:
                        tt.mul[s] = (fb[s] if c == 0 else fa[s]).data_ptr()
                shards[(c, t)] = (w, w16, accum, hyper)
                keep += [ea, es]
        nl.check(lib.zoo_eg_update(C.byref(p), st), "zoo_eg_update")
        torch.cuda.synchronize()
        for (c, t), (w, w16, accum, hyper) in shards.items():
            want = ref[(c, t)][r::W]
            torch.testing.assert_close(w[: want.shape[0]], want, atol=3e-3, rtol=3e-3)
            torch.testing.assert_close(w16[: want.shape[0]].float()[(want != full[(c, t)][r::W]).any(1)],
                                       want[(want != full[(c, t)][r::W]).any(1)], atol=2e-2, rtol=2e-2)
            assert float(accum.abs().sum()) == 0 and float(hyper[5]) == 1.0
        assert all(int(k.min()) == INT_MAX for k in reps) and all(int(k.sum()) == 0 for k in dups)     # scratch restored


@pytest.mark.parametrize("world", [1, 4])
def test_grouped_gather_matches_torch(world):
    import zoo.ops.embedding_group as G
    from zoo.native import lib as nl

    lib = G._lib()
    torch.manual_seed(1)
    B, rows_u, rows_i = 777, 403, 91
    ids = torch.stack([torch.randint(0, rows_u, (B,)), torch.randint(0, rows_i, (B,)), torch.randint(-1, rows_i, (B,))], 1).cuda()
    tabs = {"u128": torch.randn(rows_u, 128), "i64": torch.randn(rows_i, 64), "mu": torch.randn(rows_u, 32), "mi": torch.randn(rows_i, 32)}
    shard = lambda t, dt: [t[r::world].contiguous().to(dt).cuda() for r in range(world)]
    s_u, s_i = shard(tabs["u128"], torch.bfloat16), shard(tabs["i64"], torch.bfloat16)
    s_mu, s_mi = shard(tabs["mu"], torch.bfloat16), shard(tabs["mi"], torch.bfloat16)
    ld = 128 + 64 + 32
    out = torch.full((B, ld), 7.0, dtype=torch.bfloat16, device="cuda")
    sa, sb = torch.zeros(B, 32, dtype=torch.bfloat16, device="cuda"), torch.zeros(B, 32, dtype=torch.bfloat16, device="cuda")
    p = G.EgGatherParams()
    p.ids, p.id_stride, p.B, p.world, p.nspecs = ids.data_ptr(), 3, B, world, 3
    spec = [(s_u, None, rows_u, 0, 0, 0, 1, 128, 0), (s_i, None, rows_i, 0, 1, 0, 2, 64, 128), (s_mu, s_mi, rows_u, rows_i, 0, 1, 1, 32, 192)]
    for k, (ta, tb, ra, rb, c1, c2, bag, D, off) in enumerate(spec):
        e = p.s[k]
        for r in range(world):
            e.ta[r] = ta[r].data_ptr()
            if tb is not None:
                e.tb[r] = tb[r].data_ptr()
        e.out, e.ldo, e.rows_a, e.rows_b = out.data_ptr() + 2 * off, ld, ra, rb
        e.id_col, e.id_col2, e.bag, e.mean, e.D, e.dtype = c1, c2, bag, 0, D, 1
        if tb is not None:
            e.save_a, e.save_b = sa.data_ptr(), sb.data_ptr()
    nl.check(lib.zoo_eg_gather(C.byref(p), torch.cuda.current_stream().cuda_stream), "zoo_eg_gather")
    torch.cuda.synchronize()
    r16 = lambda t: t.to(torch.bfloat16).float().cuda()
    want_u = r16(tabs["u128"])[ids[:, 0]]
    ok2 = (ids[:, 2] >= 0).float()[:, None]
    want_i = r16(tabs["i64"])[ids[:, 1]] + r16(tabs["i64"])[ids[:, 2].clamp(min=0)] * ok2          # bag of 2 with a padded slot
    want_m = r16(tabs["mu"])[ids[:, 0]] * r16(tabs["mi"])[ids[:, 1]]
    torch.testing.assert_close(out[:, :128].float(), want_u, atol=1e-2, rtol=1e-2)
    torch.testing.assert_close(out[:, 128:192].float(), want_i, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(out[:, 192:].float(), want_m, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(sa.float(), r16(tabs["mu"])[ids[:, 0]])
    torch.testing.assert_close(sb.float(), r16(tabs["mi"])[ids[:, 1]])


@pytest.mark.parametrize("world", [2, 8])
def test_push_allreduce_emulated(world):
    """ar_push / ar_reduce with every rank on one GPU: all pushes first, then every rank's reduce; three steps (both mailbox
    parities are reused).  The sum of bf16-rounded copies, identical on every rank."""
    import zoo.parallel.p2p  # noqa: F401  (registers the signatures)
    from zoo.native import lib as nl

    lib = nl.cuda_lib()
    torch.manual_seed(0)
    W, n = world, 8 * 1237
    mail = [torch.zeros(2 * W * n, dtype=torch.bfloat16, device="cuda") for _ in range(W)]
    flags = [torch.zeros(2 * W * 1024, dtype=torch.int32, device="cuda") for _ in range(W)]
    epoch = [torch.zeros(2 * 1024, dtype=torch.int32, device="cuda") for _ in range(W)]
    arr = lambda ts: (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    st = torch.cuda.current_stream().cuda_stream
    grid = 5
    for step in range(3):
        g = [torch.randn(n, device="cuda") * (step + 1) for _ in range(W)]
        want = sum(x.to(torch.bfloat16).float() for x in g)
        for r in range(W):
            nl.check(lib.zoo_ar_push(g[r].data_ptr(), arr(mail), arr(flags), epoch[r].data_ptr(), n, W, r, grid, st), "zoo_ar_push")
        for r in range(W):
            nl.check(lib.zoo_ar_reduce(g[r].data_ptr(), arr(mail), arr(flags), epoch[r].data_ptr(), n, W, r, grid, st), "zoo_ar_reduce")
        torch.cuda.synchronize()
        for r in range(W):
            torch.testing.assert_close(g[r], want, atol=1e-5, rtol=1e-5)
            assert torch.equal(g[r], g[0])                  # bitwise identical replicas


def test_grouped_ncf_step_matches_per_table_path(monkeypatch):
    """The grouped path (one gather, column-fused sparse update at the end of backward) and the per-table path give the same
    NeuralCF training trajectory (SGD, so rounding noise stays at rounding size)."""
    from zoo.common.nncontext import init_nncontext
    from zoo.feature.common import MiniBatch
    from zoo.models.recommendation import NeuralCF
    from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
    from zoo.pipeline.estimator import Estimator

    ctx = init_nncontext()
    torch.manual_seed(0)
    base = NeuralCF(300, 120, 5, user_embed=64, item_embed=64, hidden_layers=[128, 64], mf_embed=32)
    g = torch.Generator().manual_seed(1)
    xs = [torch.stack([_zipf(1024, 300) + 1, _zipf(1024, 120) + 1], 1).cuda() for _ in range(6)]
    ys = [torch.randint(1, 6, (1024,), generator=g).cuda() for _ in range(6)]
    res = []
    for grouped in ("1", "0"):
        monkeypatch.setenv("ZOO_EMB_GROUPED", grouped)
        m = copy.deepcopy(base)
        est = Estimator.from_torch(model=m, optimizer=torch.optim.SGD(m.parameters(), lr=0.1),
                                   loss=SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None), ctx=ctx)
        est.engine.use_cuda_graph = False
        est.engine.prepare()
        crit = est.loss.cuda()
        losses = [float(est.engine.train_step(crit, MiniBatch(x, y, 1024))) for x, y in zip(xs, ys)]
        torch.cuda.synchronize()
        res.append((losses, {n: p.detach().clone() for n, p in m.state_dict().items()}))
    (la, sa), (lb, sb) = res
    assert max(abs(a - b) for a, b in zip(la, lb)) < 2e-3, (la, lb)
    assert la[-1] < la[0]
    for n in sa:
        if sa[n].dtype.is_floating_point:
            torch.testing.assert_close(sa[n].float(), sb[n].float(), atol=5e-3, rtol=5e-3, msg=lambda s: f"{n}: {s}")


def test_grouped_ncf_graph_matches_eager():
    from zoo.common.nncontext import init_nncontext
    from zoo.feature.common import MiniBatch
    from zoo.models.recommendation import NeuralCF
    from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
    from zoo.pipeline.estimator import Estimator

    ctx = init_nncontext()
    torch.manual_seed(0)
    base = NeuralCF(500, 300, 5, user_embed=64, item_embed=64, hidden_layers=[128, 64], mf_embed=32)
    xs = [torch.stack([torch.randint(1, 501, (1024,)), torch.randint(1, 301, (1024,))], 1).cuda() for _ in range(8)]
    ys = [torch.randint(1, 6, (1024,)).cuda() for _ in range(8)]
    out = []
    for use_graph in (False, True):
        m = copy.deepcopy(base)
        est = Estimator.from_torch(model=m, optimizer=torch.optim.SGD(m.parameters(), lr=0.1),
                                   loss=SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None), ctx=ctx)
        est.engine.use_cuda_graph = use_graph
        est.engine.prepare()
        crit = est.loss.cuda()
        losses = [float(est.engine.train_step(crit, MiniBatch(x, y, 1024))) for x, y in zip(xs, ys)]
        torch.cuda.synchronize()
        out.append((losses, m.mlp_user_embed.weight.detach().clone(), m.mf_item_embed.weight.detach().clone()))
    assert max(abs(a - b) for a, b in zip(out[0][0], out[1][0])) < 2e-3
    torch.testing.assert_close(out[0][1], out[1][1], atol=5e-3, rtol=5e-3)
    torch.testing.assert_close(out[0][2], out[1][2], atol=5e-3, rtol=5e-3)
