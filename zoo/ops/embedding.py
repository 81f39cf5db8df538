"""Embedding tables: one-sided NVLink gather forward, sort-free sparse optimizer backward.

Python face of ``csrc/ops/embedding.cu`` (kernels K2/K4).  An :class:`EmbeddingTable` owns one
row-shard of a table (``row r`` lives on rank ``r % world`` at local row ``r // world``) with
fp32 master rows, optional bf16 shadow rows for the gather, and the optimizer state of the
rows it owns.  `multi_lookup` gathers several tables into ONE activation buffer (the concat
is fused away) and `lookup_mul` produces the GMF product of two tables.

In *sparse* mode (default on GPU) the optimizer runs inside backward on the touched rows only
(lazy Adam / Adagrad / SGD, like ``torch.optim.SparseAdam``); in *dense* mode the table is an
ordinary parameter whose full gradient is produced with fp32 atomics -- that is the exact
semantics of the reference, which treats ``Embedding`` as a dense ``LookupTable`` parameter
(zoo/.../pipeline/api/keras/layers/Embedding.scala:82-101) and is what the CPU path does.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..native import lib as _nl

SPARSE_OPTS = {"sgd": 0, "adam": 1, "adagrad": 2}
_TABLE_SERIAL = [0]      # construction order is the same on every rank (sharded tables are allocated collectively)
INT_MAX = 2**31 - 1

_V, _I, _L, _F = C.c_void_p, C.c_int, C.c_longlong, C.c_float
_nl.register_cuda_sigs({
    "zoo_embedding_dispatch": [_V, _I, _I, _I, _I, _V, _L, _V, _L, _F, _L, _I, _I, _I, _V, _V, _V, _I, _V],
    "zoo_embedding_sparse_update_multi": [_V, _I, _I, _I, _I, _V, _L, _V, _L, _F, _L, _I, _I, _I, _V, _V, _V, _V, _I, _V, _V, _V, _V, _V, _I, _V],
})


class ShardRuntime:
    """Per-process state of row-sharded tables: the symmetric heap, a device-side barrier, staging buffers
    every peer can read (ids, gradient rows, GMF factors) and the duplicate scratch shared by all tables."""
    _inst = None

    @classmethod
    def get(cls):
        from ..common.nncontext import get_context

        ctx = get_context()
        if cls._inst is None or cls._inst.ctx is not ctx:
            cls._inst = cls(ctx)
        return cls._inst

    def __init__(self, ctx):
        from ..parallel.symm import get_heap, ptr_array

        if not ctx.is_gpu:
            raise _nl.NativeError("row-sharded embedding tables (world > 1) need CUDA devices")
        self.ctx, self.heap, self._ptr_array = ctx, get_heap(ctx), ptr_array
        W, r = ctx.world_size, ctx.rank
        self.flags = self.heap.alloc(4 * W)
        self.flags.tensor(torch.int32, W).zero_()
        self.epoch = torch.zeros(1, dtype=torch.int32, device=ctx.device)
        self._bflags = ptr_array(self.flags.ptrs)
        self._bepoch = ptr_array([self.epoch.data_ptr() if i == r else None for i in range(W)])
        self.stages = {}
        self.accum, self.dup = None, None
        self.pending_post = False
        torch.cuda.synchronize()
        ctx.barrier()

    def barrier(self):
        """All ranks' work queued so far on this stream is complete before anyone goes on (no host sync)."""
        _nl.check(_nl.cuda_lib().zoo_p2p_barrier(self._bflags, self._bepoch, self.ctx.world_size, self.ctx.rank, 0,
                                                 _nl.stream_ptr()), "zoo_p2p_barrier")

    def settle(self):
        """The barrier that ends a sharded backward ("every owner has finished reading my staged ids / gradient rows") is
        deferred to the next point where staged data is overwritten, so a step with several sharded lookups pays it once."""
        if self.pending_post:
            self.barrier()
            self.pending_post = False

    def stage(self, key, shape, dtype):
        """A symmetric buffer (same key -> same buffer on every rank, allocated collectively on first use)."""
        b = self.stages.get(key)
        if b is None:
            if torch.cuda.is_current_stream_capturing():
                raise _nl.NativeError("sharded-embedding staging buffers must be created before CUDA-graph capture (run one eager step first)")
            n = 1
            for d in shape:
                n *= int(d)
            buf = self.heap.alloc(max(n, 1) * torch.empty(0, dtype=dtype).element_size())
            b = self.stages[key] = (buf, buf.tensor(dtype, n).view(*shape))
        return b

    def recv_buffers(self, key, n, D):
        """Receive side of the backward dispatch for one table: ids int64 [n] (-1 = empty slot) and gradient rows bf16 [n, D]
        in the symmetric heap (peers push into them), plus this rank's per-destination counters."""
        rb = self.stages.get(("recv",) + tuple(key))
        if rb is None:
            if torch.cuda.is_current_stream_capturing():
                raise _nl.NativeError("sharded-embedding receive buffers must be created before CUDA-graph capture (run one eager step first)")
            ib, gb = self.heap.alloc(n * 8), self.heap.alloc(n * D * 2)
            rb = {"n": n, "D": D, "ids": ib.tensor(torch.int64, n), "grad": gb.tensor(torch.bfloat16, n * D).view(n, D),
                  "ids_ptrs": self._ptr_array(ib.ptrs), "grad_ptrs": self._ptr_array(gb.ptrs), "bufs": (ib, gb),
                  "counters": torch.zeros(8, dtype=torch.int32, device=self.ctx.device)}
            rb["ids"].fill_(-1)
            torch.cuda.synchronize()      # every rank's buffers are empty before any peer may push into them
            self.ctx.barrier()
            self.stages[("recv",) + tuple(key)] = rb
        return rb

    def peer_ptrs_of(self, t, byte_offset=0):
        """If tensor `t` lives inside a symmetric buffer: the addresses of the same element (+ byte_offset) in every rank's copy of
        that buffer (index = rank); else None."""
        p = t.data_ptr()
        for v in self.stages.values():
            buf = v[0] if isinstance(v, tuple) else None
            if buf is not None and buf.local_ptr <= p < buf.local_ptr + buf.nbytes:
                rel = p - buf.local_ptr + byte_offset
                return [q + rel for q in buf.ptrs]
        return None

    def scratch(self, E, D):
        if self.accum is None or self.accum.shape[0] < E or self.accum.shape[1] < D:
            E2 = max(E, 0 if self.accum is None else self.accum.shape[0])
            D2 = max(D, 0 if self.accum is None else self.accum.shape[1])
            self.accum = torch.zeros((E2, D2), dtype=torch.float32, device=self.ctx.device)
            self.dup = torch.zeros((E2,), dtype=torch.int32, device=self.ctx.device)
            self.win = torch.zeros((2 * E2 + 4,), dtype=torch.int32, device=self.ctx.device)
        return self.accum, self.dup, self.win


class EmbeddingTable(nn.Module):
    def __init__(self, num_embeddings, embedding_dim, init="uniform", init_scale=0.05, sparse=True,
                 optimizer="adam", lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
                 world=1, rank=0, shadow_bf16=None, padding_idx=None, device=None):
        super().__init__()
        self.num_embeddings, self.embedding_dim = int(num_embeddings), int(embedding_dim)
        _TABLE_SERIAL[0] += 1
        self.uid = _TABLE_SERIAL[0]      # stable key for plans / staging buffers (id() values can be reused after a free)
        self.world, self.rank = int(world), int(rank)
        self.sparse = bool(sparse)
        self.optimizer = optimizer
        self.opt_cfg = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        self.padding_idx = padding_idx
        self.rows_local = (self.num_embeddings + self.world - 1) // self.world
        self._sym = None
        if self.world > 1:
            # the shard lives in the symmetric heap so every peer can gather rows from it over NVLink
            rt = ShardRuntime.get()
            if rt.ctx.world_size != self.world or rt.ctx.rank != self.rank:
                raise ValueError(f"EmbeddingTable(world={self.world}, rank={self.rank}) does not match the context "
                                 f"(world_size={rt.ctx.world_size}, rank={rt.ctx.rank})")
            self._sym = rt.heap.alloc(self.rows_local * self.embedding_dim * 4)
            w = self._sym.tensor(torch.float32, self.rows_local * self.embedding_dim).view(self.rows_local, self.embedding_dim)
            if shadow_bf16 is None:
                # Remote gathers are NVLink-bandwidth bound and the forward rounds rows to bf16 anyway: peers read a
                # bf16 shadow of the shard (written by the sparse optimizer together with the fp32 master row).
                shadow_bf16 = os.environ.get("ZOO_EMB_SHADOW", "1") != "0"
            self._sym16 = rt.heap.alloc(self.rows_local * self.embedding_dim * 2) if shadow_bf16 else None
        else:
            w = torch.empty(self.rows_local, self.embedding_dim, device=device)  # big tables: initialise on the GPU
        if init == "uniform":
            w.uniform_(-init_scale, init_scale)
        elif init == "normal":
            w.normal_(0.0, init_scale)
        elif init == "zeros":
            w.zero_()
        else:
            nn.init.xavier_uniform_(w)
        if padding_idx is not None and self.world == 1:
            w[padding_idx].zero_()
        if shadow_bf16 is None:
            shadow_bf16 = os.environ.get("ZOO_EMB_SHADOW", "0") == "1"
        self.weight = nn.Parameter(w)
        # Sparse mode updates rows inside backward; this 1-element parameter only anchors the
        # lookup in the autograd graph (its own gradient is always None).
        self.anchor = nn.Parameter(torch.zeros(1))
        self.shadow_bf16 = shadow_bf16
        self._peer_ptrs = None if self._sym is None else list(self._sym.ptrs)   # one device pointer per rank when world > 1
        self._peer_ptrs16 = None
        if self._sym is not None and self.shadow_bf16:
            self.weight16 = self._sym16.tensor(torch.bfloat16, self.rows_local * self.embedding_dim).view(self.rows_local, self.embedding_dim)
            self.weight16.copy_(w)
            self._peer_ptrs16 = list(self._sym16.ptrs)
            self._shadow_version = self.weight._version
        if self._sym is not None:
            torch.cuda.synchronize()      # the shard is initialised before any peer may read it
            ShardRuntime.get().ctx.barrier()
        self._state_ready = False
        self._ptr_arr = None
        self.register_load_state_dict_post_hook(lambda mod, _keys: mod.refresh_shadow())

    def refresh_shadow(self):
        """Re-derive the bf16 shadow from the fp32 master rows (after a checkpoint load / in-place weight edit).  The
        sparse optimizer keeps both in step by itself; an edit through torch bumps `weight._version`, which the next
        lookup notices.  Collective for a sharded table (peers must not gather a half-refreshed shard)."""
        w16 = getattr(self, "weight16", None)
        if w16 is not None:
            with torch.no_grad():
                w16.copy_(self.weight.data)
            if self._sym is not None:
                torch.cuda.synchronize()
                ShardRuntime.get().ctx.barrier()
        self._shadow_version = self.weight._version

    def _apply(self, fn, *a, **kw):
        if self._sym is not None:
            before = self.weight.data_ptr()
        super()._apply(fn, *a, **kw)
        if self._sym is not None and self.weight.data_ptr() != before:
            raise _nl.NativeError("a row-sharded EmbeddingTable cannot be moved or cast: its shard lives in the symmetric heap")
        # On a GPU a sparse table leaves the dense optimizer: its rows are updated in backward.
        self.weight.requires_grad_(not (self.sparse and self.weight.is_cuda))
        self._state_ready = False
        return self

    @property
    def fused_sparse(self):
        return self.sparse and self.weight.is_cuda

    # ------------------------------------------------------------------ state --
    def _ensure_state(self):
        """Allocate optimizer state + backward scratch lazily on the table's device."""
        if self._state_ready and self.hyper.device == self.weight.device:
            return
        dev = self.weight.device
        z = lambda: torch.zeros_like(self.weight.data, dtype=torch.float32)
        self.exp_avg = z() if self.optimizer in ("adam", "adagrad") else None
        self.exp_avg_sq = z() if self.optimizer == "adam" else None
        if self._sym is None:
            self.weight16 = self.weight.data.to(torch.bfloat16) if self.shadow_bf16 else None
            self._shadow_version = self.weight._version
        c = self.opt_cfg
        h = torch.zeros(16, dtype=torch.float32)
        h[0], h[1], h[2], h[3], h[4], h[5], h[6] = c["lr"], c["betas"][0], c["betas"][1], c["eps"], c["weight_decay"], 0.0, 1.0
        self.hyper = h.to(dev)
        if dev.type == "cuda":
            self.rep_of_row = torch.full((self.rows_local,), INT_MAX, dtype=torch.int32, device=dev)
        self._scratch_E = 0
        self.accum = None
        self.dup_flag = None
        self._state_ready = True

    def _ensure_scratch(self, E):
        if self.world > 1:
            self.accum, self.dup_flag, self.win_list = ShardRuntime.get().scratch(E * self.world, self.embedding_dim)
            self._scratch_E = E
            return
        if self._scratch_E < E:
            dev = self.weight.device
            self.accum = torch.zeros((E, self.embedding_dim), dtype=torch.float32, device=dev)
            self.dup_flag = torch.zeros((E,), dtype=torch.int32, device=dev)
            self.win_list = torch.zeros((2 * E + 4,), dtype=torch.int32, device=dev)
            self._scratch_E = E

    def set_lr(self, lr):
        self.opt_cfg["lr"] = lr
        if self._state_ready:
            self.hyper[0] = lr

    def set_peers(self, ptrs, ptrs16=None):
        """Device pointers of every rank's shard (symmetric heap); index = rank."""
        self._peer_ptrs, self._peer_ptrs16, self._ptr_arr = list(ptrs), (list(ptrs16) if ptrs16 else None), None

    def table_ptrs(self):
        """(ctypes void*[world], dtype id) for the gather kernel."""
        use16 = self.shadow_bf16 and getattr(self, "weight16", None) is not None
        if use16 and getattr(self, "_shadow_version", None) != self.weight._version:
            self.refresh_shadow()
        if self.world == 1:
            p = [self.weight16.data_ptr() if use16 else self.weight.data_ptr()]
        else:
            p = self._peer_ptrs16 if use16 else self._peer_ptrs
            if p is None:
                raise _nl.NativeError("EmbeddingTable(world>1) needs set_peers() from the symmetric heap")
        arr = (C.c_void_p * len(p))(*p)
        return arr, (1 if use16 else 0)

    # --------------------------------------------------------------- CPU path --
    def lookup_torch(self, ids):
        return F.embedding(ids, self.weight)

    def extra_repr(self):
        return f"{self.num_embeddings}, {self.embedding_dim}, world={self.world}, sparse={self.sparse}, opt={self.optimizer}"


class _LookupSpec:
    __slots__ = ("table", "col", "bag", "mean", "mul_table", "mul_col")

    def __init__(self, table, col, bag=1, mean=False, mul_table=None, mul_col=None):
        self.table, self.col, self.bag, self.mean = table, int(col), int(bag), bool(mean)
        self.mul_table, self.mul_col = mul_table, (int(mul_col) if mul_col is not None else None)

    @property
    def dim(self):
        return self.table.embedding_dim


def _gather_into(ids, spec, out_window, save_a=None, save_b=None):
    lib = _nl.cuda_lib()
    ta, dt = spec.table.table_ptrs()
    tb, rows_b = None, 0
    if spec.mul_table is not None:
        tb, dt2 = spec.mul_table.table_ptrs()
        if dt2 != dt:
            raise _nl.NativeError("lookup_mul: both tables must use the same row dtype")
        rows_b = spec.mul_table.num_embeddings
    _nl.check(lib.zoo_embedding_gather(ids.data_ptr(), ids.stride(0), spec.col, spec.mul_col or 0, ids.shape[0], spec.bag,
                                       int(spec.mean), ta, spec.table.num_embeddings, tb, rows_b, spec.table.world,
                                       spec.dim, dt, out_window.data_ptr(), out_window.stride(0), _nl.ptr(save_a),
                                       _nl.ptr(save_b), _nl.stream_ptr()), "zoo_embedding_gather")


def _sparse_update_sharded(table, ids_buf, ids, col, bag, dout_buf, dout_off, ldg, mul_buf, grad_scale):
    """Backward of a row-sharded table: this rank pulls, from every peer's staged ids / gradient rows, the
    entries whose rows it owns (the all-to-all happens inside the kernels)."""
    lib, rt = _nl.cuda_lib(), ShardRuntime.get()
    table._ensure_state()
    B = ids.shape[0]
    table._ensure_scratch(B * bag)
    pa = rt._ptr_array
    slot = ids.numel() * 8
    ids_p = pa([ids_buf.local_ptr + s * slot for s in range(table.world)])     # every rank's ids, replicated locally in forward
    dout_p = pa([p + dout_off * 2 for p in dout_buf.ptrs])
    mul_p = pa(mul_buf.ptrs) if mul_buf is not None else None
    D = table.embedding_dim
    accum = table.accum if table.accum.shape[1] == D else table.accum.view(-1)[: B * bag * table.world * D].view(-1, D)
    _nl.check(lib.zoo_embedding_sparse_update_multi(
        ids_p, ids.stride(0), col, B, bag, dout_p, ldg, mul_p, D if mul_buf is not None else 0, float(grad_scale),
        table.num_embeddings, table.world, table.rank, D, table.weight.data_ptr(), _nl.ptr(getattr(table, "weight16", None)), _nl.ptr(table.exp_avg),
        _nl.ptr(table.exp_avg_sq), SPARSE_OPTS[table.optimizer], table.hyper.data_ptr(), table.rep_of_row.data_ptr(),
        accum.data_ptr(), table.dup_flag.data_ptr(), table.win_list.data_ptr(), 1, _nl.stream_ptr()), "zoo_embedding_sparse_update")


def _push_mode():
    """Backward of row-sharded tables: "pull" (default; owners read ids / gradient rows of their entries from the peers'
    staging buffers inside the update kernels) or "push" (sources dispatch gradient rows into the owners' receive buffers,
    owners then update from local memory).  Measured on 4 / 8 B200s: within 2-3 % of each other (profiles/README.md)."""
    return os.environ.get("ZOO_SHARD_BWD", "pull") == "push"


def _dispatch(table, rb, ids, col, bag, dout_window, mul, grad_scale):
    """Push this rank's gradient rows for `table` into the owners' receive buffers (zoo_embedding_dispatch)."""
    B = ids.shape[0]
    _nl.check(_nl.cuda_lib().zoo_embedding_dispatch(
        ids.data_ptr(), ids.stride(0), col, B, bag, dout_window.data_ptr(), dout_window.stride(0), _nl.ptr(mul),
        mul.stride(0) if mul is not None else 0, float(grad_scale), table.num_embeddings, table.world, table.rank,
        table.embedding_dim, rb["ids_ptrs"], rb["grad_ptrs"], rb["counters"].data_ptr(), B * bag, _nl.stream_ptr()), "zoo_embedding_dispatch")


def _sparse_update_received(table, rb):
    """Owner side of the dispatch: the ordinary single-GPU sparse update over the receive buffer (local row ids, final
    gradient rows), then the buffer's id slots are emptied for the next step."""
    lib = _nl.cuda_lib()
    table._ensure_state()
    n, D = rb["n"], table.embedding_dim
    table._ensure_scratch(n // table.world)
    accum = table.accum if table.accum.shape[1] == D else table.accum.view(-1)[: n * D].view(-1, D)
    _nl.check(lib.zoo_embedding_sparse_update_w(
        rb["ids"].data_ptr(), 1, 0, n, 1, rb["grad"].data_ptr(), D, None, 0, 1.0, table.rows_local, 1, 0, D,
        table.weight.data_ptr(), _nl.ptr(getattr(table, "weight16", None)), _nl.ptr(table.exp_avg), _nl.ptr(table.exp_avg_sq),
        SPARSE_OPTS[table.optimizer], table.hyper.data_ptr(), table.rep_of_row.data_ptr(), accum.data_ptr(),
        table.dup_flag.data_ptr(), table.win_list.data_ptr(), 1, _nl.stream_ptr()), "zoo_embedding_sparse_update_w")
    rb["ids"].fill_(-1)


def _sparse_update(table, ids, col, bag, dout_window, mul, grad_scale):
    lib = _nl.cuda_lib()
    table._ensure_state()
    E = ids.shape[0] * bag
    table._ensure_scratch(E)
    _nl.check(lib.zoo_embedding_sparse_update_w(
        ids.data_ptr(), ids.stride(0), col, ids.shape[0], bag, dout_window.data_ptr(), dout_window.stride(0),
        _nl.ptr(mul), mul.stride(0) if mul is not None else 0, float(grad_scale), table.num_embeddings, table.world,
        table.rank, table.embedding_dim, table.weight.data_ptr(), _nl.ptr(getattr(table, "weight16", None)),
        _nl.ptr(table.exp_avg), _nl.ptr(table.exp_avg_sq), SPARSE_OPTS[table.optimizer], table.hyper.data_ptr(),
        table.rep_of_row.data_ptr(), table.accum.data_ptr(), table.dup_flag.data_ptr(), table.win_list.data_ptr(), 1,
        _nl.stream_ptr()), "zoo_embedding_sparse_update_w")


def _dense_grad(table, ids, col, bag, dout_window, mul, grad_scale):
    lib = _nl.cuda_lib()
    g = getattr(table.weight, "_zoo_grad", None)
    ret = None
    if g is None:
        g = ret = torch.zeros_like(table.weight, dtype=torch.float32)
    _nl.check(lib.zoo_embedding_dense_grad(ids.data_ptr(), ids.stride(0), col, ids.shape[0], bag, dout_window.data_ptr(),
                                           dout_window.stride(0), _nl.ptr(mul), mul.stride(0) if mul is not None else 0,
                                           float(grad_scale), table.num_embeddings, table.embedding_dim, g.data_ptr(),
                                           _nl.stream_ptr()), "zoo_embedding_dense_grad")
    return ret


class _MultiLookupFn(torch.autograd.Function):
    """Gather every spec into its column window of one [B, sum(D)] bf16 buffer."""
    offered_dx_out = None

    @staticmethod
    def forward(ctx, ids, specs, *weights):
        B = ids.shape[0]
        total = sum(s.dim for s in specs)
        ld = (total + 7) // 8 * 8
        out = torch.empty((B, ld), dtype=torch.bfloat16, device=ids.device)
        if ld != total:
            out[:, total:].zero_()
        sharded = any(t.world > 1 for t in _flat_tables(specs))
        ctx.sharded = sharded
        ctx.push = sharded and _push_mode()
        key = None
        if ctx.push:
            # backward DISPATCHES gradient rows to their owners (zoo_embedding_dispatch): forward stages nothing
            rt = ShardRuntime.get()
            rt.settle()
            key = ("lookup", tuple((s.table.uid, s.col, s.bag, s.mul_table.uid if s.mul_table is not None else 0, s.mul_col) for s in specs), B, ids.shape[1])
            if torch.is_grad_enabled():
                for k, s in enumerate(specs):      # receive buffers exist (collectively allocated) before the first backward
                    for which, t in enumerate([s.table] + ([s.mul_table] if s.mul_table is not None else [])):
                        if t.world > 1:
                            rt.recv_buffers((key, k, which), t.world * B * s.bag, t.embedding_dim)
        elif sharded:
            # ids (and GMF factors below) are staged in symmetric buffers: peers read them in backward
            rt = ShardRuntime.get()
            rt.settle()        # peers are done with what the previous backward staged before anything is staged again
            key = ("lookup", tuple((s.table.uid, s.col, s.bag, s.mul_table.uid if s.mul_table is not None else 0, s.mul_col) for s in specs), B, ids.shape[1])
            # every rank PUSHES its id matrix into slot [rank] of every peer's copy, so the owner-side passes of backward
            # (election, duplicate accumulation) scan all world * B entries from local memory
            W, r = rt.ctx.world_size, rt.ctx.rank
            ids_buf, ids_all = rt.stage(key + ("ids",), (W,) + tuple(ids.shape), torch.int64)
            src = ids.contiguous()
            _nl.check(_nl.cuda_lib().zoo_symm_bcast(rt._ptr_array(ids_buf.ptrs), r * src.numel() * 8, src.data_ptr(), src.numel() * 8, W,
                                                    _nl.stream_ptr()), "zoo_symm_bcast")
            ids = ids_all[r]
            ctx.ids_buf = ids_buf
        saves, off, ctx.mul_bufs = [], 0, []
        for k, s in enumerate(specs):
            s.table._ensure_state()
            sa = sb = None
            if s.mul_table is not None:
                s.mul_table._ensure_state()
                if sharded and not ctx.push:
                    ba, sa = rt.stage(key + ("sa", k), (B, s.dim), torch.bfloat16)
                    bb, sb = rt.stage(key + ("sb", k), (B, s.dim), torch.bfloat16)
                    ctx.mul_bufs.append((ba, bb))
                else:
                    sa = torch.empty((B, s.dim), dtype=torch.bfloat16, device=ids.device)
                    sb = torch.empty((B, s.dim), dtype=torch.bfloat16, device=ids.device)
                    ctx.mul_bufs.append(None)
            else:
                ctx.mul_bufs.append(None)
            _gather_into(ids, s, out[:, off:off + s.dim], sa, sb)
            saves += [sa, sb]
            off += s.dim
        ctx.specs = specs
        ctx.total = total
        ctx.key = key
        ctx.save_for_backward(ids, *[t for t in saves if t is not None])
        ctx.has_save = [t is not None for t in saves]
        res = out[:, :total]
        if sharded and not ctx.push and torch.is_grad_enabled():
            # the gradient rows of this lookup must sit where peers can read them: offer the staging buffer to the consumer
            # (FusedDense writes its input gradient straight into it), which saves the copy in backward
            dbuf, dst = rt.stage(key + ("dout",), (B, ld), torch.bfloat16)
            _MultiLookupFn.offered_dx_out = dst[:, :total]      # attached to the result by multi_lookup()
        return res

    @staticmethod
    def backward(ctx, dout):
        ids = ctx.saved_tensors[0]
        saved = list(ctx.saved_tensors[1:])
        specs = ctx.specs
        d16 = dout if dout.dtype == torch.bfloat16 else dout.to(torch.bfloat16)
        if d16.stride(-1) != 1:
            d16 = d16.contiguous()
        rt = dbuf = None
        jobs = []
        if ctx.push:
            rt = ShardRuntime.get()
        elif ctx.sharded:
            # stage the gradient rows where peers can read them; the barrier makes every rank's staging
            # (and every rank's forward gathers of the old weights) complete before rows are updated
            rt = ShardRuntime.get()
            B = ids.shape[0]
            ld = (ctx.total + 7) // 8 * 8
            dbuf, dst = rt.stage(ctx.key + ("dout",), (B, ld), torch.bfloat16)
            if d16.data_ptr() != dst.data_ptr() or d16.stride(0) != dst.stride(0):
                dst[:, :ctx.total].copy_(d16)
            rt.barrier()
        grads, off, si = [], 0, 0
        for k, s in enumerate(specs):
            win = d16[:, off:off + s.dim]
            sa = sb = None
            if ctx.has_save[2 * k]:
                sa, sb = saved[si], saved[si + 1]
                si += 2
            scale = 1.0
            if s.mean and s.bag > 1:
                # mean pooling divides by the number of PRESENT ids of the bag (the reference's sparse `mean` combiner; the forward
                # kernel does the same), so every sample's gradient row is scaled by 1 / count before it is scattered
                idc = ids[:, s.col:s.col + s.bag]
                cnt = ((idc >= 0) & (idc < s.table.num_embeddings)).sum(1).clamp(min=1).to(torch.float32)
                win = (win.float() / cnt[:, None]).to(torch.bfloat16)
            pairs = [(s.table, s.col, sb, 1)]
            if s.mul_table is not None:
                pairs.append((s.mul_table, s.mul_col, sa, 0))
            for table, col, mul, which in pairs:
                if table.world > 1 and ctx.push:
                    if not table.sparse:
                        raise _nl.NativeError("row-sharded tables are trained with the sparse optimizer (sparse=True)")
                    rb = rt.recv_buffers((ctx.key, k, 0 if which == 1 else 1), table.world * ids.shape[0] * s.bag, table.embedding_dim)
                    # 1/world: the loss is a mean over the GLOBAL batch, like the data-parallel dense part
                    _dispatch(table, rb, ids, col, s.bag, win, mul, scale / table.world)
                    jobs.append((table, rb))
                    grads.append(None)
                elif table.world > 1:
                    if s.mean and s.bag > 1:
                        raise NotImplementedError("mean-pooled bags over row-sharded tables are not supported (use sum pooling)")
                    if not table.sparse:
                        raise _nl.NativeError("row-sharded tables are trained with the sparse optimizer (sparse=True)")
                    mb = ctx.mul_bufs[k][which] if (mul is not None and ctx.mul_bufs[k] is not None) else None
                    # 1/world: the loss is a mean over the GLOBAL batch, like the data-parallel dense part
                    _sparse_update_sharded(table, ctx.ids_buf, ids, col, s.bag, dbuf, off, (ctx.total + 7) // 8 * 8, mb, scale / table.world)
                    grads.append(None)
                elif table.sparse:
                    _sparse_update(table, ids, col, s.bag, win, mul, scale)
                    grads.append(None)
                else:
                    grads.append(_dense_grad(table, ids, col, s.bag, win, mul, scale))
            off += s.dim
        if jobs:
            # every rank's pushes have landed (and every rank's forward gathers of the old rows are done) before owners update
            rt.barrier()
            for table, rb in jobs:
                _sparse_update_received(table, rb)
        if ctx.sharded:
            rt.pending_post = True     # nobody overwrites staging buffers / gathers new rows before all owners are done: see settle()
        return (None, None, *grads)


def _flat_tables(specs):
    ts = []
    for s in specs:
        ts.append(s.table)
        if s.mul_table is not None:
            ts.append(s.mul_table)
    return ts


def multi_lookup(ids: torch.Tensor, specs, anchor: torch.Tensor | None = None, splits=None, embed=None):
    """ids: int64 [B, ncols].  specs: list of _LookupSpec.  Returns bf16 [B, sum(D)] on GPU, fp32 on CPU; with `splits`
    (a list of column widths summing to sum(D)) a tuple of those column windows instead."""
    if splits is not None:
        if ids.dim() == 1:
            ids = ids.unsqueeze(1)
        if ids.is_cuda:
            if ids.dtype != torch.int64:
                ids = ids.long()
            from . import embedding_group as _eg

            if len(specs) <= _eg.MAX_SPECS and _eg.eligible(specs):
                # embed = (split index, width, column): that split is produced inside a fresh [B, width] buffer, which is
                # returned as one more element of the tuple (see zoo.ops.embedding_group.join_columns)
                return _eg.grouped_lookup(ids, specs, splits, embed)
        full = multi_lookup(ids, specs, anchor)
        outs, off = [], 0
        for w in splits:
            outs.append(full[:, off:off + w])
            off += w
        if embed is not None:
            outs.append(None)      # no fused buffer on this path: the caller concatenates
        return tuple(outs)
    if ids.dim() == 1:
        ids = ids.unsqueeze(1)
    if ids.is_cuda:
        if ids.dtype != torch.int64:
            ids = ids.long()
        from . import embedding_group as _eg

        if len(specs) <= _eg.MAX_SPECS and _eg.eligible(specs):
            # every lookup in one gather launch; sparse updates of all tables at the end of backward (column groups)
            return _eg.grouped_lookup(ids, specs)
        weights = [t.anchor if t.sparse else t.weight for t in _flat_tables(specs)]
        _MultiLookupFn.offered_dx_out = None
        res = _MultiLookupFn.apply(ids, specs, *weights)
        if _MultiLookupFn.offered_dx_out is not None:
            res._zoo_dx_out = _MultiLookupFn.offered_dx_out
            _MultiLookupFn.offered_dx_out = None
        return res
    outs = []
    for s in specs:
        idc = ids[:, s.col:s.col + s.bag].long()
        ok = (idc >= 0) & (idc < s.table.num_embeddings)      # out-of-range ids contribute nothing, as in the kernel
        e = s.table.lookup_torch(idc.clamp(0, s.table.num_embeddings - 1)) * ok.unsqueeze(-1)  # [B, bag, D]
        if s.mul_table is not None:
            idb = ids[:, s.mul_col:s.mul_col + s.bag].long()
            okb = (idb >= 0) & (idb < s.mul_table.num_embeddings)
            e = e * (s.mul_table.lookup_torch(idb.clamp(0, s.mul_table.num_embeddings - 1)) * okb.unsqueeze(-1))
        e = e.sum(1) / ok.sum(1, keepdim=True).clamp(min=1).to(e.dtype) if (s.mean and s.bag > 1) else e.sum(1)
        outs.append(e)
    return torch.cat(outs, dim=1) if len(outs) > 1 else outs[0]


def lookup(ids, table, col=0, bag=1, mean=False, anchor=None):
    return multi_lookup(ids, [_LookupSpec(table, col, bag, mean)], anchor)


def lookup_mul(ids, table_a, col_a, table_b, col_b, anchor=None):
    """GMF branch: table_a[ids[:, col_a]] * table_b[ids[:, col_b]]."""
    return multi_lookup(ids, [_LookupSpec(table_a, col_a, 1, False, table_b, col_b)], anchor)


LookupSpec = _LookupSpec
