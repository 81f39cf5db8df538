"""Fused reduce-scatter + optimizer + all-gather kernel (K1) and the flag barrier.

Single-GPU tests run all ranks inside ONE cooperative launch (`emulate=1`): kernels that wait on
one another must never be separate launches on one GPU.  The multi-GPU test spawns one process per
GPU with torchrun and is skipped on boxes with fewer than 2 GPUs.
"""
import ctypes as C
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


def _arr(ts):
    return (C.c_void_p * len(ts))(*[t.data_ptr() if t is not None else None for t in ts])


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("opt", ["adam", "sgd"])
@pytest.mark.parametrize("replicated", [0, 1])
def test_rs_opt_ag_emulated_matches_torch(world, opt, replicated):
    import zoo.parallel.symm  # registers signatures
    from zoo.native import lib as nl
    from zoo.parallel.flat import OPT_IDS, OptConfig

    lib = nl.cuda_lib()
    torch.manual_seed(0)
    n = 4 * world * 1000
    p0 = torch.randn(n, device="cuda")
    ps = [p0.clone() for _ in range(world)]
    p16 = [p0.to(torch.bfloat16) for _ in range(world)]
    gs = [torch.randn(n, device="cuda") for _ in range(world)]
    ms = [torch.zeros(n, device="cuda") for _ in range(world)]
    vs = [torch.zeros(n, device="cuda") for _ in range(world)]
    flags = [torch.zeros(2 * world * 1024, dtype=torch.int32, device="cuda") for _ in range(world)]
    epochs = [torch.zeros(1024, dtype=torch.int32, device="cuda") for _ in range(world)]
    cfg = OptConfig("adam", lr=1e-2) if opt == "adam" else OptConfig("sgd", lr=1e-2, b1=0.9)
    ref_p = p0.clone().requires_grad_(True)
    ref_opt = torch.optim.Adam([ref_p], lr=1e-2) if opt == "adam" else torch.optim.SGD([ref_p], lr=1e-2, momentum=0.9)
    st = torch.cuda.current_stream().cuda_stream
    for step in range(1, 4):
        hv = torch.tensor(cfg.hyper_vector(step=float(step), gscale=1.0 / world), dtype=torch.float32, device="cuda")
        hy = [hv] * world
        gsum = sum(gs) / world
        nl.check(lib.zoo_rs_opt_ag(_arr(gs), _arr(ps), _arr(p16), _arr(ms), _arr(vs), _arr(flags), _arr(epochs), _arr(hy),
                                   0, n, world, 0, 1, OPT_IDS[cfg.name], 1 | (2 * replicated), 8, st), "zoo_rs_opt_ag")
        ref_p.grad = gsum.clone()
        ref_opt.step()
        torch.cuda.synchronize()
        for r in range(world):
            torch.testing.assert_close(ps[r], ref_p.detach(), atol=1e-5, rtol=1e-4)
            torch.testing.assert_close(p16[r].float(), ref_p.detach(), atol=2e-2, rtol=2e-2)
            assert float(gs[r].abs().sum()) == 0.0       # gradients zeroed for the next iteration
        gs = [g.normal_() for g in gs]
    assert int(epochs[0][0]) == 3


def test_p2p_barrier_emulated():
    import zoo.parallel.symm  # noqa: F401
    from zoo.native import lib as nl

    lib = nl.cuda_lib()
    world = 8
    flags = [torch.zeros(world, dtype=torch.int32, device="cuda") for _ in range(world)]
    epochs = [torch.zeros(1, dtype=torch.int32, device="cuda") for _ in range(world)]
    for _ in range(5):
        nl.check(lib.zoo_p2p_barrier(_arr(flags), _arr(epochs), world, 0, 1, torch.cuda.current_stream().cuda_stream), "bar")
    torch.cuda.synchronize()
    assert all(int(e[0]) == 5 for e in epochs) and all(int(f.min()) == 5 for f in flags)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_dp_p2p_matches_nccl_multi_gpu():
    n = min(torch.cuda.device_count(), 8)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "tests", "mgpu_dp_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "MGPU_DP_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_sharded_embedding_matches_single_gpu():
    n = 2
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29534", os.path.join(root, "tests", "mgpu_sharded_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "MGPU_SHARDED_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("with_mul", [False, True])
def test_sparse_update_multi_source_on_one_gpu(world, with_mul):
    """The pull-based sharded backward, with all 'ranks' on one GPU (their launches do not wait on each other):
    every rank scans all ranks' entries and updates the rows it owns; the union must equal one un-sharded update."""
    import zoo.ops.embedding as E
    from zoo.native import lib as nl

    lib = nl.cuda_lib()
    torch.manual_seed(0)
    rows, D, B = 101, 16, 64
    ids = [torch.randint(0, rows, (B, 2), device="cuda") for _ in range(world)]
    dout = [torch.randn(B, D, device="cuda").to(torch.bfloat16) for _ in range(world)]
    mul = [torch.randn(B, D, device="cuda").to(torch.bfloat16) for _ in range(world)] if with_mul else None
    w_full = torch.randn(rows, D, device="cuda")
    hyper = torch.zeros(16, device="cuda"); hyper[0], hyper[1], hyper[2], hyper[3], hyper[5] = 0.1, 0.9, 0.999, 1e-8, 1.0
    # reference: dense gradient + one Adam step on touched rows
    g = torch.zeros(rows, D, device="cuda")
    for s in range(world):
        gr = dout[s].float() * (mul[s].float() if with_mul else 1.0)
        g.index_add_(0, ids[s][:, 1], gr)
    touched = g.abs().sum(1) > 0
    m = 0.1 * g; v = 0.001 * g * g
    ref = w_full - 0.1 * (m / 0.1) / ((v / 0.001).sqrt() + 1e-8)
    ref = torch.where(touched[:, None], ref, w_full)
    st = torch.cuda.current_stream().cuda_stream
    arr = lambda ts: (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    for r in range(world):
        rl = (rows + world - 1) // world
        w = torch.zeros(rl, D, device="cuda")
        w[: w_full[r::world].shape[0]] = w_full[r::world]
        ea, es = torch.zeros_like(w), torch.zeros_like(w)
        rep = torch.full((rl,), 2**31 - 1, dtype=torch.int32, device="cuda")
        accum = torch.zeros(world * B, D, device="cuda"); dup = torch.zeros(world * B, dtype=torch.int32, device="cuda")
        win = torch.full((2 * world * B + 4,), -7, dtype=torch.int32, device="cuda")
        nl.check(lib.zoo_embedding_sparse_update_multi(arr(ids), 2, 1, B, 1, arr(dout), D, arr(mul) if with_mul else None, D, 1.0, rows, world, r, D,
                                                       w.data_ptr(), None, ea.data_ptr(), es.data_ptr(), 1, hyper.data_ptr(), rep.data_ptr(),
                                                       accum.data_ptr(), dup.data_ptr(), win.data_ptr(), 0, st), "multi")
        torch.cuda.synchronize()
        want = ref[r::world]
        torch.testing.assert_close(w[: want.shape[0]], want, atol=2e-3, rtol=2e-3)
        assert int(rep.min()) == 2**31 - 1 and float(accum.abs().sum()) == 0 and int(dup.sum()) == 0     # scratch restored


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("with_mul", [False, True])
def test_dispatch_then_owner_update_on_one_gpu(world, with_mul):
    """Push-based backward of a sharded table, all ranks emulated on one GPU: every source dispatches its gradient rows
    into the owners' receive buffers, then each owner runs the single-GPU sparse update over what it received."""
    import ctypes as C

    import zoo.ops.embedding  # registers signatures
    from zoo.native import lib as nl

    lib = nl.cuda_lib()
    torch.manual_seed(5)
    rows, D, B = 1003, 64, 700
    arr = lambda ts: (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    w_full = torch.randn(rows, D, device="cuda")
    ids = [torch.stack([torch.zeros(B, dtype=torch.int64), torch.randint(0, rows, (B,))], 1).cuda() for _ in range(world)]
    dout = [torch.randn(B, D, device="cuda").to(torch.bfloat16) for _ in range(world)]
    mul = [torch.randn(B, D, device="cuda").to(torch.bfloat16) for _ in range(world)]
    # reference: dense gradient (bf16-rounded rows like the transfer), one Adam step on touched rows
    g = torch.zeros(rows, D, device="cuda")
    for s in range(world):
        row = dout[s].float() * (mul[s].float() if with_mul else 1.0) * 0.5
        g.index_add_(0, ids[s][:, 1], row.to(torch.bfloat16).float())
    touched = torch.zeros(rows, dtype=torch.bool, device="cuda")
    for s in range(world):
        touched[ids[s][:, 1]] = True
    m = 0.1 * g
    v = 0.001 * g * g
    ref = w_full - 1e-2 * (m / 0.1) / ((v / 0.001).sqrt() + 1e-8) * touched.unsqueeze(1)
    cap = B
    recv_ids = [torch.full((world * cap,), -1, dtype=torch.int64, device="cuda") for _ in range(world)]
    recv_grad = [torch.zeros(world * cap, D, dtype=torch.bfloat16, device="cuda") for _ in range(world)]
    st = torch.cuda.current_stream().cuda_stream
    for s in range(world):          # every source pushes
        counters = torch.zeros(8, dtype=torch.int32, device="cuda")
        nl.check(lib.zoo_embedding_dispatch(ids[s].data_ptr(), 2, 1, B, 1, dout[s].data_ptr(), D, mul[s].data_ptr() if with_mul else None, D, 0.5,
                                            rows, world, s, D, arr(recv_ids), arr(recv_grad), counters.data_ptr(), cap, st), "dispatch")
        torch.cuda.synchronize()
        assert int(counters[:world].sum()) == B
    for r in range(world):          # every owner updates its shard from what it received
        hyper = torch.tensor([1e-2, 0.9, 0.999, 1e-8, 0.0, 0.0, 1.0] + [0.0] * 9, device="cuda")   # step 0 -> bumped to 1
        rl = (rows + world - 1) // world
        w = torch.zeros(rl, D, device="cuda")
        w[: w_full[r::world].shape[0]] = w_full[r::world]
        ea, es = torch.zeros_like(w), torch.zeros_like(w)
        rep = torch.full((rl,), 2**31 - 1, dtype=torch.int32, device="cuda")
        n = world * cap
        accum = torch.zeros(n, D, device="cuda"); dup = torch.zeros(n, dtype=torch.int32, device="cuda")
        win = torch.zeros(2 * n + 4, dtype=torch.int32, device="cuda")
        nl.check(lib.zoo_embedding_sparse_update_w(recv_ids[r].data_ptr(), 1, 0, n, 1, recv_grad[r].data_ptr(), D, None, 0, 1.0, rl, 1, 0, D,
                                                   w.data_ptr(), None, ea.data_ptr(), es.data_ptr(), 1, hyper.data_ptr(), rep.data_ptr(),
                                                   accum.data_ptr(), dup.data_ptr(), win.data_ptr(), 1, st), "update")
        torch.cuda.synchronize()
        want = ref[r::world]
        torch.testing.assert_close(w[: want.shape[0]], want, atol=2e-3, rtol=2e-3)
        assert int(rep.min()) == 2**31 - 1 and float(accum.abs().sum()) == 0 and int(dup.sum()) == 0
