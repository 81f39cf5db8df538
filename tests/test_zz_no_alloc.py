"""Allocation-free sparse embedding update, called through the C ABI.

Some users of libzoo_cuda.so apply embedding updates from places where they may not allocate
device memory (a serving thread that must never hit cudaMalloc's device-wide synchronisation, a
host callback inside another runtime's stream, a process whose memory is fully committed to a
pool).  `zoo_embedding_sparse_update_noalloc` serves them with scratch that lives in a static
device array inside the library.  This test drives it with ctypes exactly as a C caller would
and checks that two consecutive calls give the same table and optimizer state as the regular
entry point fed with caller-owned scratch -- the second call also proves that the shared array
is re-initialised correctly after it has been used.  The file name makes pytest collect it last.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

INT_MAX = 2**31 - 1
ARGS_COMMON = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong,
               C.c_float, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
               C.c_int, C.c_void_p]


def _lib():
    from zoo.native.lib import cuda_lib_path

    lib = C.CDLL(str(cuda_lib_path()))
    lib.zoo_embedding_sparse_update.argtypes = ARGS_COMMON + [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.zoo_embedding_sparse_update.restype = C.c_int
    lib.zoo_embedding_sparse_update_noalloc.argtypes = ARGS_COMMON + [C.c_void_p]
    lib.zoo_embedding_sparse_update_noalloc.restype = C.c_int
    lib.zoo_last_error.restype = C.c_char_p
    return lib


def _hyper(step):
    return torch.tensor([0.05, 0.9, 0.999, 1e-8, 0.0, float(step), 1.0] + [0.0] * 9, dtype=torch.float32, device="cuda")


@pytest.mark.parametrize("opt", [1, 2])  # Adam, Adagrad
def test_noalloc_matches_regular_entry_point_twice(opt):
    lib = _lib()
    torch.manual_seed(0)
    rows, D, B = 5000, 64, 4096          # many duplicate ids: exercises the shared accumulation rows
    stream = torch.cuda.current_stream().cuda_stream
    w0 = torch.randn(rows, D, device="cuda") * 0.1
    state = {}
    for name in ("ref", "new"):
        state[name] = dict(w=w0.clone(), m=torch.zeros(rows, D, device="cuda"), v=torch.zeros(rows, D, device="cuda"))
    rep = torch.full((rows,), INT_MAX, dtype=torch.int32, device="cuda")
    accum = torch.zeros(B, D, device="cuda")
    dup = torch.zeros(B, dtype=torch.int32, device="cuda")
    for call in (1, 2):
        ids = torch.randint(0, rows, (B, 1), device="cuda")
        dout = torch.randn(B, D, device="cuda").to(torch.bfloat16)
        hyper = _hyper(call)
        common = lambda s: [ids.data_ptr(), 1, 0, B, 1, dout.data_ptr(), D, None, 0, 1.0, rows, 1, 0, D, s["w"].data_ptr(),
                            None, s["m"].data_ptr(), s["v"].data_ptr(), opt, hyper.data_ptr()]
        rc = lib.zoo_embedding_sparse_update(*common(state["ref"]), rep.data_ptr(), accum.data_ptr(), dup.data_ptr(), stream)
        assert rc == 0, lib.zoo_last_error()
        rc = lib.zoo_embedding_sparse_update_noalloc(*common(state["new"]), stream)
        assert rc == 0, lib.zoo_last_error()
        torch.cuda.synchronize()
        # same tolerance as tests/test_gpu_ops.py::test_sparse_update_matches_dense_reference
        torch.testing.assert_close(state["new"]["w"], state["ref"]["w"], atol=2e-3, rtol=2e-2)
        torch.testing.assert_close(state["new"]["m"], state["ref"]["m"], atol=2e-3, rtol=2e-2)
        if opt == 1:
            torch.testing.assert_close(state["new"]["v"], state["ref"]["v"], atol=2e-3, rtol=2e-2)
    assert not torch.equal(state["new"]["w"], w0)


def test_noalloc_rejects_oversized_problems():
    lib = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    ids = torch.zeros(1, 1, dtype=torch.int64, device="cuda")
    dout = torch.zeros(1, 256, dtype=torch.bfloat16, device="cuda")
    w = torch.zeros(4, 256, device="cuda")
    hyper = _hyper(1)
    base = [ids.data_ptr(), 1, 0, 1, 1, dout.data_ptr(), 256, None, 0, 1.0]
    tail = [w.data_ptr(), None, w.data_ptr(), w.data_ptr(), 0, hyper.data_ptr(), stream]
    assert lib.zoo_embedding_sparse_update_noalloc(*base, 4, 1, 0, 256, *tail) == 3            # D > 128
    assert b"exceeds the static scratch" in lib.zoo_last_error()
    assert lib.zoo_embedding_sparse_update_noalloc(*base, (1 << 20) + 1, 1, 0, 64, *tail) == 3  # too many rows
    big = [ids.data_ptr(), 1, 0, 8193, 1, dout.data_ptr(), 256, None, 0, 1.0]
    assert lib.zoo_embedding_sparse_update_noalloc(*big, 4, 1, 0, 64, *tail) == 3               # too many entries


def test_noalloc_on_a_shard_of_a_two_rank_table():
    """world = 2: the caller owns the odd rows of a table with an odd row count (local rows = rows / world rounded up); only
    entries of its own rows are applied, exactly like the regular entry point."""
    lib = _lib()
    torch.manual_seed(1)
    rows, D, B, world, rank = 5001, 32, 2048, 2, 1
    rows_local = (rows + world - 1) // world
    stream = torch.cuda.current_stream().cuda_stream
    w0 = torch.randn(rows_local, D, device="cuda") * 0.1
    st = {n: dict(w=w0.clone(), m=torch.zeros(rows_local, D, device="cuda"), v=torch.zeros(rows_local, D, device="cuda")) for n in ("ref", "new")}
    rep = torch.full((rows_local,), INT_MAX, dtype=torch.int32, device="cuda")
    accum = torch.zeros(B, D, device="cuda")
    dup = torch.zeros(B, dtype=torch.int32, device="cuda")
    for call in (1, 2):
        ids = torch.randint(0, rows, (B, 1), device="cuda")
        dout = torch.randn(B, D, device="cuda").to(torch.bfloat16)
        hyper = _hyper(call)
        common = lambda s: [ids.data_ptr(), 1, 0, B, 1, dout.data_ptr(), D, None, 0, 1.0, rows, world, rank, D, s["w"].data_ptr(),
                            None, s["m"].data_ptr(), s["v"].data_ptr(), 1, hyper.data_ptr()]
        assert lib.zoo_embedding_sparse_update(*common(st["ref"]), rep.data_ptr(), accum.data_ptr(), dup.data_ptr(), stream) == 0
        assert lib.zoo_embedding_sparse_update_noalloc(*common(st["new"]), stream) == 0, lib.zoo_last_error()
        torch.cuda.synchronize()
        torch.testing.assert_close(st["new"]["w"], st["ref"]["w"], atol=2e-3, rtol=2e-2)
        torch.testing.assert_close(st["new"]["v"], st["ref"]["v"], atol=2e-3, rtol=2e-2)
    owned = torch.zeros(rows_local, dtype=torch.bool, device="cuda")
    owned[(ids[:, 0][ids[:, 0] % world == rank]) // world] = True
    assert not torch.equal(st["new"]["w"][owned], w0[owned])
