"""Grouped embedding path: every lookup of a step in ONE gather launch, one routing launch per step for row-sharded tables,
and the sparse backward of all tables in three launches at the END of backward.

Python face of ``csrc/ops/embedding_group.cu`` (parameter blocks in ``include/zoo_emb_group.h``).  What it changes against the
per-table path of :mod:`zoo.ops.embedding`:

* tables that are looked up with the same id column (NeuralCF: MLP-user + MF-user, MLP-item + MF-item) form a *column group*
  with one election of representatives, one winner list and one apply launch for all of their rows;
* the sparse updates are not launched from the individual backward nodes: each node records what it needs and the first one
  queues a callback on the autograd engine, which runs when backward is complete and then issues ONE device barrier (when tables
  are sharded) and the three update launches for every group of the step;
* with row-sharded tables every source routes its entries to the owners in forward (``zoo_eg_route``), so the owner's backward
  passes run over the ~B entries it owns; gradient rows stay where the producing kernel wrote them (symmetric staging memory)
  and are read in place over NVLink by the owner's accumulate / apply kernels.

Reference semantics: zoo Embedding's LookupTable forward / accGradParameters + the optimizer step on the touched rows
(zoo/.../pipeline/api/keras/layers/Embedding.scala:82-101, keras/optimizers/Adam.scala:59-106).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from ..native import lib as _nl

MAX_PEERS, MAX_GROUPS, MAX_TABS, MAX_SPECS = 8, 4, 2, 4
_V, _I, _L, _F = C.c_void_p, C.c_int, C.c_longlong, C.c_float


class EgTab(C.Structure):
    _fields_ = [("w", _V), ("w16", _V), ("m", _V), ("v", _V), ("accum", _V), ("hyper", _V), ("dout", _V * MAX_PEERS),
                ("mul", _V * MAX_PEERS), ("ldg", _L), ("ldmul", _L), ("D", _I), ("opt", _I), ("grad_scale", _F), ("bump_step", _I)]


class EgGroup(C.Structure):
    _fields_ = [("ids", _V), ("pairs", _V), ("counts", _V), ("rep_of_row", _V), ("dup_flag", _V), ("win_list", _V),
                ("win_count", _V), ("rows", _L), ("id_stride", _I), ("id_col", _I), ("bag", _I), ("E_local", _I), ("cap", _I),
                ("world", _I), ("rank", _I), ("ntabs", _I), ("max_entries", _I), ("pad_", _I), ("t", EgTab * MAX_TABS)]


class EgParams(C.Structure):
    _fields_ = [("ngroups", _I), ("pad_", _I), ("g", EgGroup * MAX_GROUPS)]


class EgSpec(C.Structure):
    _fields_ = [("ta", _V * MAX_PEERS), ("tb", _V * MAX_PEERS), ("out", _V), ("save_a", _V), ("save_b", _V), ("rows_a", _L),
                ("rows_b", _L), ("ldo", _L), ("id_col", _I), ("id_col2", _I), ("bag", _I), ("mean", _I), ("D", _I), ("dtype", _I)]


class EgGatherParams(C.Structure):
    _fields_ = [("ids", _V), ("id_stride", _I), ("B", _I), ("world", _I), ("nspecs", _I), ("s", EgSpec * MAX_SPECS)]


class EgRouteGroup(C.Structure):
    _fields_ = [("pairs", _V * MAX_PEERS), ("counts", _V * MAX_PEERS), ("counters", _V), ("rows", _L), ("id_col", _I), ("bag", _I)]


class EgRouteParams(C.Structure):
    _fields_ = [("ids", _V), ("id_stride", _I), ("B", _I), ("world", _I), ("rank", _I), ("cap", _I), ("ngroups", _I),
                ("g", EgRouteGroup * MAX_GROUPS)]


_nl.register_cuda_sigs({
    "zoo_eg_gather": [_V, _V],
    "zoo_eg_route": [_V, _V],
    "zoo_eg_update": [_V, _V],
    "zoo_eg_sizes": [_V],
})
_checked = False
SPARSE_OPTS = {"sgd": 0, "adam": 1, "adagrad": 2}


def _lib():
    global _checked
    lib = _nl.cuda_lib()
    if not _checked:
        sizes = (C.c_int * 6)()
        if lib.zoo_eg_sizes(sizes) != 0 or list(sizes) != [C.sizeof(t) for t in (EgTab, EgGroup, EgParams, EgSpec, EgGatherParams, EgRouteParams)]:
            raise _nl.NativeError(f"zoo_emb_group.h and its ctypes mirror disagree: C {list(sizes)} vs python "
                                  f"{[C.sizeof(t) for t in (EgTab, EgGroup, EgParams, EgSpec, EgGatherParams, EgRouteParams)]}")
        _checked = True
    return lib


def enabled():
    return os.environ.get("ZOO_EMB_GROUPED", "1") != "0"


def eligible(specs):
    """The grouped kernels cover fused-sparse tables with 16-byte rows; everything else keeps the per-table path."""
    if not enabled() or not specs:
        return False
    worlds = set()
    for s in specs:
        for t in ([s.table] + ([s.mul_table] if s.mul_table is not None else [])):
            if not t.fused_sparse or t.embedding_dim % 8 or t.embedding_dim > 256 or t.optimizer not in SPARSE_OPTS:
                return False
            worlds.add(t.world)
        if s.mean and s.bag > 1:
            return False
        if s.mul_table is not None and (s.bag != 1 or s.mul_table.embedding_dim != s.table.embedding_dim):
            return False
    return len(worlds) == 1


def _shard_ptrs(table):
    arr, dt = table.table_ptrs()
    return [arr[i] for i in range(len(arr))], dt


class _Plan:
    """Static description of one grouped lookup (same specs, batch size and id layout): parameter blocks, scratch, and -- for
    sharded tables -- the mailboxes and staging buffers in the symmetric heap.  Built once, before any CUDA-graph capture."""

    def __init__(self, specs, B, id_stride, device):
        from .embedding import ShardRuntime

        self.specs, self.B, self.id_stride, self.device = list(specs), int(B), int(id_stride), device
        t0 = specs[0].table
        self.world, self.rank = t0.world, t0.rank
        self.sharded = self.world > 1
        self.rt = ShardRuntime.get() if self.sharded else None
        self.total = sum(s.dim for s in specs)
        self.ld = (self.total + 7) // 8 * 8
        self.offsets, off = [], 0
        for s in specs:
            self.offsets.append(off)
            off += s.dim
        if torch.cuda.is_current_stream_capturing():
            raise _nl.NativeError("an embedding lookup plan must be built before CUDA-graph capture (run one eager step first)")
        # ---- column groups: tables sharing (id column, bag) ------------------------------------------------------------------
        groups = {}
        for k, s in enumerate(specs):
            groups.setdefault((s.col, s.bag), []).append((s.table, k, "a"))
            if s.mul_table is not None:
                groups.setdefault((s.mul_col, s.bag), []).append((s.mul_table, k, "b"))
        self.groups = []
        for (col, bag), members in groups.items():
            for i in range(0, len(members), MAX_TABS):
                self.groups.append({"col": col, "bag": bag, "members": members[i:i + MAX_TABS]})
        # ---- staging (sharded): gradient rows and product factors where peers can read them ----------------------------------
        self.key = ("egplan", tuple((s.table.uid, s.col, s.bag, s.mul_table.uid if s.mul_table is not None else 0, s.mul_col) for s in specs), self.B,
                self.id_stride)
        self.dout_buf = self.dout = None
        self.saves = {}
        for k, s in enumerate(specs):
            if s.mul_table is None:
                continue
            if self.sharded:
                ba, sa = self.rt.stage(self.key + ("sa", k), (self.B, s.dim), torch.bfloat16)
                bb, sb = self.rt.stage(self.key + ("sb", k), (self.B, s.dim), torch.bfloat16)
                self.saves[k] = (sa, sb, ba, bb)
            else:
                self.saves[k] = None       # allocated per call (they are saved tensors of the autograd node)
        if self.sharded:
            self.dout_buf, self.dout = self.rt.stage(self.key + ("dout",), (self.B, self.ld), torch.bfloat16)
        # ---- scratch per group -------------------------------------------------------------------------------------------------
        W = self.world
        for g in self.groups:
            E = self.B * g["bag"]
            g["E"] = E
            g["cap"] = E
            g["maxE"] = E if W == 1 else W * E
            tabs = [m[0] for m in g["members"]]
            for t in tabs:
                t._ensure_state()
            g["dup_flag"] = torch.zeros(MAX_TABS * g["maxE"], dtype=torch.int32, device=device)
            g["win_list"] = torch.zeros(2 * g["maxE"] + 2, dtype=torch.int32, device=device)
            g["win_count"] = torch.zeros(2, dtype=torch.int32, device=device)
            g["accum"] = [torch.zeros((g["maxE"], t.embedding_dim), dtype=torch.float32, device=device) for t in tabs]
        if self.sharded:
            # one mailbox per routed (column, bag, row count): groups that only differ in their tables share it
            self._route_groups, seen = [], {}
            for g in self.groups:
                rk = (g["col"], g["bag"], g["members"][0][0].num_embeddings)
                if rk in seen:
                    g["pairs_buf"], g["counts_buf"] = seen[rk]["pairs_buf"], seen[rk]["counts_buf"]
                    continue
                g["pairs_buf"] = self.rt.heap.alloc(W * g["cap"] * 8)
                g["counts_buf"] = self.rt.heap.alloc(4 * MAX_PEERS)
                g["counters"] = torch.zeros(MAX_PEERS + 1, dtype=torch.int32, device=device)
                seen[rk] = g
                self._route_groups.append(g)
            if len(self._route_groups) > MAX_GROUPS:
                raise _nl.NativeError(f"a grouped lookup routes at most {MAX_GROUPS} id columns; split the lookup")
            torch.cuda.synchronize()
            self.rt.ctx.barrier()       # every rank's mailboxes exist (and are zero) before any peer routes into them
        self._gather = None
        self._route = None
        self.keep = None

    # ---- parameter blocks -------------------------------------------------------------------------------------------------------
    def gather_params(self, ids, out, saves):
        p = self._gather
        if p is None:
            p = self._gather = EgGatherParams()
            p.id_stride, p.B, p.world, p.nspecs = ids.stride(0), self.B, self.world, len(self.specs)
            for k, s in enumerate(self.specs):
                e = p.s[k]
                pa, dt = _shard_ptrs(s.table)
                for r, ptr in enumerate(pa):
                    e.ta[r] = ptr
                if s.mul_table is not None:
                    pb, dt2 = _shard_ptrs(s.mul_table)
                    if dt2 != dt:
                        raise _nl.NativeError("lookup_mul: both tables must use the same row dtype")
                    for r, ptr in enumerate(pb):
                        e.tb[r] = ptr
                    e.rows_b, e.id_col2 = s.mul_table.num_embeddings, s.mul_col
                e.rows_a, e.id_col, e.bag, e.mean, e.D, e.dtype = s.table.num_embeddings, s.col, s.bag, int(s.mean), s.dim, dt
        else:
            for s in self.specs:       # shadow rows may have been re-derived (checkpoint load): keep them in step
                s.table.table_ptrs()
                if s.mul_table is not None:
                    s.mul_table.table_ptrs()
        p.ids = ids.data_ptr()
        p.id_stride = ids.stride(0)
        for k, s in enumerate(self.specs):
            e = p.s[k]
            dst, rel = out[k]          # (buffer, column offset inside it) for this lookup
            e.out = dst.data_ptr() + 2 * rel
            e.ldo = dst.stride(0)
            sv = saves.get(k)
            e.save_a, e.save_b = (sv[0].data_ptr(), sv[1].data_ptr()) if sv is not None else (None, None)
        return p

    def route_params(self, ids):
        p = self._route
        if p is None:
            p = self._route = EgRouteParams()
            p.B, p.world, p.rank, p.ngroups = self.B, self.world, self.rank, len(self.route_groups())
            p.cap = max(g["cap"] for g in self.route_groups())
            for k, g in enumerate(self.route_groups()):
                e = p.g[k]
                for r in range(self.world):
                    e.pairs[r] = g["pairs_buf"].ptrs[r]
                    e.counts[r] = g["counts_buf"].ptrs[r]
                e.counters = g["counters"].data_ptr()
                e.rows = g["members"][0][0].num_embeddings
                e.id_col, e.bag = g["col"], g["bag"]
        p.ids = ids.data_ptr()
        p.id_stride = ids.stride(0)
        return p

    def route_groups(self):
        return self._route_groups


_PLANS = {}
_PENDING = []
PRE_FLUSH_HOOKS = []      # callables run right before the sparse updates of a step (the engine pushes dense gradients here)


def _plan_for(specs, ids):
    key = (tuple((s.table.uid, s.col, s.bag, s.mean, s.mul_table.uid if s.mul_table is not None else 0, s.mul_col) for s in specs), ids.shape[0],
            ids.shape[1], ids.device)
    p = _PLANS.get(key)
    if p is None:
        p = _PLANS[key] = _Plan(specs, ids.shape[0], ids.shape[1], ids.device)
        p.keep = list(specs)       # id()-keyed cache: keep the keyed objects alive
    return p


def scratch_is_clean():
    """True when every plan's backward scratch is back at its start values (duplicate flags and accumulation rows zero)."""
    for p in _PLANS.values():
        for g in p.groups:
            if int(g["dup_flag"].sum()) != 0 or any(float(a.abs().sum()) != 0.0 for a in g["accum"]):
                return False
    return True


def reset_plans():
    _PLANS.clear()
    _PENDING.clear()


class _GroupedLookupFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, plan, splits, training, embed, *anchors):
        lib = _lib()
        B = ids.shape[0]
        if plan.sharded:
            plan.rt.settle()       # owners are done with what the previous step routed / staged before anything is overwritten
            _nl.check(lib.zoo_eg_route(C.byref(plan.route_params(ids)), _nl.stream_ptr()), "zoo_eg_route")
        out = torch.empty((B, plan.ld), dtype=torch.bfloat16, device=ids.device)
        if plan.ld != plan.total:
            out[:, plan.total:].zero_()
        dests = [(out, plan.offsets[k]) for k in range(len(plan.specs))]
        wide = None
        if embed is not None:
            # one split lands directly in a column window of a wider, fresh buffer (the consumer's concat is fused away): the
            # other columns of that buffer are filled by whoever produces them (FusedDense(out=...)), see `join_columns`
            j, width, col = embed
            wide = torch.empty((B, width), dtype=torch.bfloat16, device=ids.device)
            lo = sum(splits[:j])
            for k in range(len(plan.specs)):
                if lo <= plan.offsets[k] < lo + splits[j]:
                    dests[k] = (wide, col + plan.offsets[k] - lo)
        saves = {}
        for k, s in enumerate(plan.specs):
            if s.mul_table is None:
                continue
            if plan.sharded:
                saves[k] = plan.saves[k][:2]
            elif training:       # (grad mode is off inside Function.forward: the caller passes it in)
                saves[k] = (torch.empty((B, s.dim), dtype=torch.bfloat16, device=ids.device),
                            torch.empty((B, s.dim), dtype=torch.bfloat16, device=ids.device))
        _nl.check(lib.zoo_eg_gather(C.byref(plan.gather_params(ids, dests, saves)), _nl.stream_ptr()), "zoo_eg_gather")
        ctx.plan, ctx.splits = plan, splits
        ctx.saves = saves
        flat = [ids] + [t for k in sorted(saves) for t in saves[k]]
        ctx.save_for_backward(*flat)
        if splits is None:
            return out[:, :plan.total]
        outs, off = [], 0
        for j, w in enumerate(splits):
            if embed is not None and j == embed[0]:
                outs.append(wide[:, embed[2]:embed[2] + w])
            else:
                outs.append(out[:, off:off + w])
            off += w
        if wide is not None:
            ctx.mark_non_differentiable(wide)
            ctx.set_materialize_grads(False)
            outs.append(wide)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *douts):
        plan = ctx.plan
        ids = ctx.saved_tensors[0]
        saved = list(ctx.saved_tensors[1:])
        saves, si = {}, 0
        for k in sorted(ctx.saves):
            saves[k] = (saved[si], saved[si + 1])
            si += 2
        # windows: (column offset, width, gradient tensor) per forward output
        widths = [plan.total] if ctx.splits is None else list(ctx.splits)
        wins, off = [], 0
        for w, d in zip(widths, douts):
            if d is not None:
                d16 = d if d.dtype == torch.bfloat16 else d.to(torch.bfloat16)
                if d16.stride(-1) != 1 or d16.stride(0) % 4 or d16.data_ptr() % 8:
                    d16 = d16.contiguous()
                wins.append((off, w, d16))
            off += w
        _PENDING.append({"plan": plan, "ids": ids, "wins": wins, "saves": saves})
        if len(_PENDING) == 1:
            torch.autograd.Variable._execution_engine.queue_callback(flush_pending)
        return (None,) * len(ctx.needs_input_grad)


def _window_for(wins, col_off, width):
    """The output-gradient tensor that covers columns [col_off, col_off + width) of the lookup's buffer, and the offset of that
    range inside it."""
    for off, w, d in wins:
        if off <= col_off and col_off + width <= off + w:
            return d, col_off - off
    return None, 0


def flush_pending():
    """End of backward: one barrier (sharded tables), then elect / accumulate / apply for every column group of the step."""
    if not _PENDING:
        return
    jobs, _PENDING[:] = list(_PENDING), []
    lib = _lib()
    st = _nl.stream_ptr()
    sharded = [j for j in jobs if j["plan"].sharded]
    for hook in list(PRE_FLUSH_HOOKS):
        hook()
    blocks, keep = [], []
    for j in jobs:
        plan, ids = j["plan"], j["ids"]
        rt = plan.rt
        for g in plan.groups:
            e = EgGroup()
            tabs = g["members"]
            e.ids, e.id_stride, e.id_col, e.bag, e.E_local = ids.data_ptr(), ids.stride(0), g["col"], g["bag"], g["E"]
            e.world, e.rank, e.ntabs, e.cap, e.max_entries = plan.world, plan.rank, len(tabs), g["cap"], g["maxE"]
            e.rows = tabs[0][0].num_embeddings
            for table, _, _ in tabs:
                table._ensure_state()
            e.rep_of_row, e.dup_flag = tabs[0][0].rep_of_row.data_ptr(), g["dup_flag"].data_ptr()
            e.win_list, e.win_count = g["win_list"].data_ptr(), g["win_count"].data_ptr()
            if plan.sharded:
                e.pairs, e.counts = g["pairs_buf"].local_ptr, g["counts_buf"].local_ptr
            skip = False
            for q, (table, k, which) in enumerate(tabs):
                s = plan.specs[k]
                d, rel = _window_for(j["wins"], plan.offsets[k], s.dim)
                if d is None:
                    skip = True        # this lookup's output received no gradient
                    break
                t = e.t[q]
                t.w, t.w16 = table.weight.data_ptr(), _nl.ptr(getattr(table, "weight16", None))
                t.m, t.v = _nl.ptr(table.exp_avg), _nl.ptr(table.exp_avg_sq)
                t.accum, t.hyper = g["accum"][q].data_ptr(), table.hyper.data_ptr()
                t.D, t.opt, t.bump_step = table.embedding_dim, SPARSE_OPTS[table.optimizer], 1
                # 1/world: the loss is a mean over the GLOBAL batch, like the data-parallel dense part
                t.grad_scale = 1.0 / plan.world
                mul = None
                if s.mul_table is not None:
                    sa, sb = j["saves"][k]
                    mul = sb if which == "a" else sa
                if plan.sharded:
                    ptrs = rt.peer_ptrs_of(d, rel * 2)
                    if ptrs is None:       # the producer did not write into peer-readable memory: stage the window
                        dst = plan.dout[:, plan.offsets[k]:plan.offsets[k] + s.dim]
                        dst.copy_(d[:, rel:rel + s.dim])
                        ptrs = [p + 2 * plan.offsets[k] for p in plan.dout_buf.ptrs]
                        t.ldg = plan.dout.stride(0)
                    else:
                        t.ldg = d.stride(0)
                    for r in range(plan.world):
                        t.dout[r] = ptrs[r]
                    if mul is not None:
                        mb = plan.saves[k][3] if which == "a" else plan.saves[k][2]
                        for r in range(plan.world):
                            t.mul[r] = mb.ptrs[r]
                        t.ldmul = mul.stride(0)
                else:
                    t.dout[0] = d.data_ptr() + 2 * rel
                    t.ldg = d.stride(0)
                    if mul is not None:
                        t.mul[0] = mul.data_ptr()
                        t.ldmul = mul.stride(0)
                keep.append((d, mul))
            if not skip:
                blocks.append((e, {id(tb[0]) for tb in tabs}))
    if sharded:
        # every rank's gradient rows are staged and every rank's forward gathers of the old rows are done before owners update
        sharded[0]["plan"].rt.barrier()
    # groups of one launch run concurrently: a table (and its row map) appears in at most one group per launch
    chunks, cur, used = [], [], set()
    for e, tids in blocks:
        if len(cur) == MAX_GROUPS or (used & tids):
            chunks.append(cur)
            cur, used = [], set()
        cur.append(e)
        used |= tids
    if cur:
        chunks.append(cur)
    for chunk in chunks:
        p = EgParams()
        p.ngroups = len(chunk)
        for k, e in enumerate(chunk):
            p.g[k] = e
        _nl.check(lib.zoo_eg_update(C.byref(p), st), "zoo_eg_update")
        _nl.LAUNCH_COUNT["n"] += 2          # three kernels per call
    if sharded:
        sharded[0]["plan"].rt.pending_post = True   # settle(): nobody routes / stages / gathers again before all owners are done
    del keep


class _JoinFn(torch.autograd.Function):
    """`wide` already holds its column windows (each written in place by its producer); this node only ties the producers'
    autograd histories to the joined tensor.  Backward hands every producer its column window of the gradient (views)."""

    @staticmethod
    def forward(ctx, wide, *parts):
        ctx.widths = [p.shape[1] for p in parts]
        return wide.view_as(wide)

    @staticmethod
    def backward(ctx, d):
        outs, off = [], 0
        for w in ctx.widths:
            outs.append(d[:, off:off + w])
            off += w
        return (None, *outs)


def join_columns(wide, *parts):
    """torch.cat(parts, 1) for parts that were produced directly inside `wide` (in column order)."""
    return _JoinFn.apply(wide, *parts)


def grouped_lookup(ids, specs, splits=None, embed=None):
    """ids int64 [B, ncols] on a CUDA device; returns bf16 [B, sum(D)] or, with `splits` (column widths), a tuple of windows of
    that buffer.  In a sharded run the attribute `_zoo_dx_out` of each result is peer-readable memory its gradient may be written
    into directly (FusedDense does)."""
    plan = _plan_for(specs, ids)
    anchors = []
    for s in specs:
        anchors.append(s.table.anchor)
        if s.mul_table is not None:
            anchors.append(s.mul_table.anchor)
    res = _GroupedLookupFn.apply(ids, plan, tuple(splits) if splits else None, torch.is_grad_enabled(), embed, *anchors)
    if plan.sharded and torch.is_grad_enabled():
        if splits is None:
            res._zoo_dx_out = plan.dout[:, :plan.total]
        else:
            off = 0
            for r, w in zip(res, splits):
                r._zoo_dx_out = plan.dout[:, off:off + w]
                off += w
    return res
