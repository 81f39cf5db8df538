"""P2PFusedSync: the product path of data-parallel training on GPUs.

Gradient reduce-scatter, optimizer and weight all-gather are ONE kernel over NVLink peer memory
(`csrc/comm/rs_opt_ag.cu`, kernel K1); this module owns its plumbing: the three arenas live in the
symmetric heap, flags and epochs are allocated once, and a step is `zoo_hyper_step` +
`zoo_rs_opt_ag` per parameter group.  torch.distributed (NCCL) is used only to exchange IPC
handles and to broadcast the initial weights.
"""
from __future__ import annotations

import os

import torch

import ctypes as C

from ..native import lib as _nl
from .flat import OPT_IDS
from .symm import get_heap, ptr_array

_nl.register_cuda_sigs({
    "zoo_ar_push": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_void_p],
    "zoo_rs_opt_ag_ex": [C.c_void_p] * 8 + [C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                         C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p],
    "zoo_ar_reduce": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_void_p],
})
AR_MAX_CTAS = 1024      # kArMaxCtas of csrc/comm/ar_push.cu


class P2PFusedSync:
    name = "p2p"

    def __init__(self, ctx):
        if not ctx.is_gpu:
            raise RuntimeError("P2P path needs CUDA devices")
        self.ctx = ctx
        self.heap = get_heap(ctx)
        self.fopt = None
        self._bufs = []
        self.flag_stride = 1024      # kFlagStride of csrc/comm/rs_opt_ag.cu
        self.max_grid = 592          # 4 CTAs per SM: a ResNet-sized arena needs the whole GPU to fill NVLink (64 CTAs reached 11 % of it)

    # -- arena allocation in the symmetric heap -------------------------------------------
    def allocator(self):
        def alloc(n, dtype):
            esize = torch.empty(0, dtype=dtype).element_size()
            # arenas of the owner-computes kernel are also bound to an NVLS multicast object (in-switch gradient reduction and
            # one-store weight broadcast); small arenas take the push all-reduce and need none
            buf = self.heap.alloc(n * esize, multicast=not self._is_small(n))
            self._bufs.append(buf)
            return buf.tensor(dtype, n)

        return alloc

    def _is_small(self, n):
        mode = os.environ.get("ZOO_K1_SMALL", "auto")
        return (mode == "1") or (mode == "auto" and n * 2 * self.ctx.world_size <= (64 << 20))

    def pad_to(self):
        import math

        return math.lcm(8, 4 * self.ctx.world_size)      # ar_push moves 8-float units, rs_opt_ag float4 units per rank

    def graph_safe(self):
        f = getattr(self, "fopt", None)
        if f is None or getattr(self, "small", False):
            return True                 # push all-reduce + local fused optimizer (clipping included): kernels only
        return f.clip_l2 is None      # large arenas with L2 clipping take the library all-reduce: no capture

    def attach(self, fopt):
        self.fopt = fopt
        W = self.ctx.world_size
        fopt.set_grad_scale(1.0 / W)
        a = fopt.arena
        by_ptr = {b.local_ptr: b for b in self._bufs}
        self.buf_p, self.buf_g = by_ptr[a.flat_p.data_ptr()], by_ptr[a.flat_g.data_ptr()]
        self.buf_p16 = by_ptr[a.flat_p16.data_ptr()] if a.flat_p16 is not None else None
        self.flag_buf = self.heap.alloc(4 * 2 * W * self.flag_stride)
        self.epoch = torch.zeros(self.flag_stride, dtype=torch.int32, device=self.ctx.device)
        self.bar_flags = self.heap.alloc(4 * W)
        self.bar_epoch = torch.zeros(1, dtype=torch.int32, device=self.ctx.device)
        # flags count epochs from zero: clear them on every rank before any peer can signal
        self.flag_buf.tensor(torch.int32, 2 * W * self.flag_stride).zero_()
        self.bar_flags.tensor(torch.int32, W).zero_()
        torch.cuda.synchronize()
        self.ctx.barrier()
        r = self.ctx.rank
        self._g = ptr_array(self.buf_g.ptrs)
        self._p = ptr_array(self.buf_p.ptrs)
        self._p16 = ptr_array(self.buf_p16.ptrs) if self.buf_p16 is not None else None
        loc = lambda t: ptr_array([t.data_ptr() if (t is not None and i == r) else None for i in range(W)])
        self._m, self._v = loc(fopt.m), loc(fopt.v)
        self._flags = ptr_array(self.flag_buf.ptrs)
        self._epoch = loc(self.epoch)
        self._bflags = ptr_array(self.bar_flags.ptrs)
        self._bepoch = loc(self.bar_epoch)
        # Small arenas (a recommender's dense part): one-handshake push all-reduce with a bf16 wire (csrc/comm/ar_push.cu),
        # then the ordinary local fused optimizer -- clipping included -- on the summed gradient.
        n = a.numel
        self.small = self._is_small(n)
        # NVLS (in-switch reduction / broadcast): measured on 2 GPUs the multicast path is SLOWER than peer loads and stores
        # (0.420 vs 0.302 ms for 25.6 M parameters: with two ranks the switch saves nothing and every access, the local copy
        # included, crosses NVLink); it pays when one load replaces many, so `auto` takes it from 4 ranks on.  ZOO_NVLS=1 / 0 force it.
        want = os.environ.get("ZOO_NVLS", "auto")
        self.nvls = ((not self.small) and self.buf_g.mc_ptr is not None and self.buf_p.mc_ptr is not None
                     and (want == "1" or (want == "auto" and W >= 4)))
        self._pushed = False
        if self.small:
            if n % 8:
                raise RuntimeError("flat arena length must be a multiple of 8 (pad_to)")
            self.mail = self.heap.alloc(2 * W * n * 2)
            self.ar_flags = self.heap.alloc(4 * 2 * W * AR_MAX_CTAS)
            self.ar_flags.tensor(torch.int32, 2 * W * AR_MAX_CTAS).zero_()
            self.ar_epoch = torch.zeros(2 * AR_MAX_CTAS, dtype=torch.int32, device=self.ctx.device)
            self._mail, self._arflags = ptr_array(self.mail.ptrs), ptr_array(self.ar_flags.ptrs)
            self.ar_grid = int(min(592, max(1, (n // 8 + 255) // 256)))
            torch.cuda.synchronize()
            self.ctx.barrier()
            from ..ops import embedding_group as _eg

            # dense gradients are final when backward ends: push them before the sparse embedding updates start, so the
            # flags have long arrived when the reduce runs after those updates
            import weakref

            ref = weakref.ref(self)

            def _hook():
                me = ref()
                if me is None:              # the engine that owned this exchange is gone
                    _eg.PRE_FLUSH_HOOKS.remove(_hook)
                elif me.fopt is not None and me._armed:
                    me.push_gradients()

            self._hook = _hook
            self._armed = True
            _eg.PRE_FLUSH_HOOKS.append(_hook)

    def broadcast_parameters(self):
        import torch.distributed as dist

        a = self.fopt.arena
        dist.broadcast(a.flat_p, src=0)
        if a.flat_p16 is not None:
            a.flat_p16.copy_(a.flat_p)
        torch.cuda.synchronize()
        dist.barrier()

    def barrier(self):
        """Device-side barrier on the current stream (no host sync, no NCCL)."""
        _nl.check(_nl.cuda_lib().zoo_p2p_barrier(self._bflags, self._bepoch, self.ctx.world_size, self.ctx.rank, 0,
                                                 _nl.stream_ptr()), "zoo_p2p_barrier")

    def _grid_for(self, n, replicated=False):
        # every thread's loads cross NVLink (~2-3 us each): a range of a few hundred KB wants ONE float4 per thread so all
        # of them are in flight at once; large ranges settle at the CTA cap and stream
        per_rank_units = n // 4 if replicated else n // (4 * self.ctx.world_size)
        return int(min(self.max_grid, max(4, (per_rank_units + 255) // 256)))

    def _replicated(self, n):
        """Small ranges: every rank sums all gradient copies and updates its own replica (one NVLink round trip less)."""
        mode = os.environ.get("ZOO_K1_REPLICATED", "auto")
        if mode in ("0", "1"):
            return mode == "1"
        return n * 4 * self.ctx.world_size <= (32 << 20)

    def push_gradients(self):
        """First half of the small-arena exchange (no waiting): bf16 copies of this rank's gradients into every mailbox."""
        if not self.small or self._pushed:
            return
        a = self.fopt.arena
        _nl.check(_nl.cuda_lib().zoo_ar_push(a.flat_g.data_ptr(), self._mail, self._arflags, self.ar_epoch.data_ptr(), a.numel,
                                            self.ctx.world_size, self.ctx.rank, self.ar_grid, _nl.stream_ptr()), "zoo_ar_push")
        self._pushed = True

    def begin_step(self):
        """Small arenas finish with the local fused optimizer (`f.step()`): its hyper-parameter refresh can start now, on a side
        stream.  The owner-computes path of large arenas refreshes inside its own step."""
        if self.small and self.fopt.clip_l2 is None:
            self.fopt.begin_step()

    def step(self):
        f = self.fopt
        if self.small:
            self.push_gradients()
            self._pushed = False
            a = f.arena
            _nl.check(_nl.cuda_lib().zoo_ar_reduce(a.flat_g.data_ptr(), self._mail, self._arflags, self.ar_epoch.data_ptr(), a.numel,
                                                  self.ctx.world_size, self.ctx.rank, self.ar_grid, _nl.stream_ptr()), "zoo_ar_reduce")
            f.step()        # clip (on the averaged gradient) + optimizer + bf16 shadows + zero the gradients, all local kernels
            return
        if f.clip_l2 is not None:
            # the global gradient norm is only known after the whole reduction, which the owner-computes kernel does not have:
            # large arenas with L2 clipping take the library all-reduce + local fused optimizer (same arenas, same state)
            import torch.distributed as dist

            dist.all_reduce(f.arena.flat_g)
            f.step()
            return
        f.step_count += 1
        f._apply_schedules()
        refresh = f._sync_hyper()
        lib, st, W, r = _nl.cuda_lib(), _nl.stream_ptr(), self.ctx.world_size, self.ctx.rank
        for i, g in enumerate(f.groups):
            hp = f.hyper_dev[i].data_ptr()
            _nl.check(lib.zoo_hyper_step(hp, f.hyper_stage[i].data_ptr() if refresh else None, st), "zoo_hyper_step")
            hy = ptr_array([hp if k == r else None for k in range(W)])
            rep = self._replicated(g.end - g.start)
            mc = lambda b: b.mc_ptr if (self.nvls and b is not None and b.mc_ptr) else None      # (the kernel adds the range offset)
            clip = f.clip_const is not None
            lo, hi = ((f.clip_const[0] / f.grad_scale, f.clip_const[1] / f.grad_scale) if clip else (0.0, 0.0))
            _nl.check(lib.zoo_rs_opt_ag_ex(self._g, self._p, self._p16, self._m, self._v, self._flags, self._epoch, hy,
                                           g.start, g.end, W, r, 0, OPT_IDS[g.cfg.name], 1 | (2 if rep else 0),
                                           self._grid_for(g.end - g.start, rep), mc(self.buf_g), mc(self.buf_p),
                                           mc(self.buf_p16), int(clip), lo, hi, st), "zoo_rs_opt_ag")

    def gather_optimizer_state(self):
        """Owner-computes ranges keep m / v only on the rank that owns a shard; before a checkpoint every rank's copy is made
        complete (collective, NCCL -- checkpoint time, not the step)."""
        import torch.distributed as dist

        f, W, r = self.fopt, self.ctx.world_size, self.ctx.rank
        if self.small:
            return          # every rank runs the full optimizer on the summed gradient: state is complete everywhere
        for g in f.groups:
            n = g.end - g.start
            if self._replicated(n):
                continue
            shard = n // W
            for t in (f.m, f.v):
                if t is None:
                    continue
                full = t[g.start:g.end]
                own = full[r * shard:(r + 1) * shard].clone()
                dist.all_gather_into_tensor(full, own)
        torch.cuda.synchronize()

    def close(self):
        h = getattr(self, "_hook", None)
        if h is not None:
            from ..ops import embedding_group as _eg

            if h in _eg.PRE_FLUSH_HOOKS:
                _eg.PRE_FLUSH_HOOKS.remove(h)
            self._hook = None
