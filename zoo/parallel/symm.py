"""Symmetric heap: device buffers every rank can address directly (component C1).

Each rank `cudaMalloc`s a slab, exports it with CUDA IPC, and maps every peer's slab; after this
bootstrap (one `all_gather_object` over torch.distributed) kernels read and write peer memory
through plain pointers over NVLink -- the transport that replaces the reference's Spark
BlockManager puts/fetches (Topology.scala:1128-1206) and Ray's object store
(pyzoo/zoo/ray/raycontext.py:118-140).
"""
from __future__ import annotations

import ctypes as C

import torch

from ..native import lib as _nl

_nl.register_cuda_sigs({
    "zoo_symm_alloc": [C.c_longlong, C.c_void_p, C.c_void_p],
    "zoo_symm_open": [C.c_void_p, C.c_void_p],
    "zoo_symm_close": [C.c_void_p],
    "zoo_symm_free": [C.c_void_p],
    "zoo_rs_opt_ag": [C.c_void_p] * 8 + [C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p],
    "zoo_p2p_barrier": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p],
    "zoo_symm_bcast": [C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p],
    "zoo_vmm_alloc": [C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p],
    "zoo_vmm_import": [C.c_int, C.c_longlong, C.c_void_p],
    "zoo_vmm_free": [C.c_void_p, C.c_longlong],
    "zoo_close_fd": [C.c_int],
    "zoo_mc_supported": [C.c_void_p],
    "zoo_mc_granularity": [C.c_int, C.c_void_p],
    "zoo_vmm_alloc_aligned": [C.c_longlong, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p],
    "zoo_mc_create": [C.c_longlong, C.c_int, C.c_void_p, C.c_void_p],
    "zoo_mc_import": [C.c_int, C.c_void_p],
    "zoo_mc_add_device": [C.c_ulonglong],
    "zoo_mc_bind": [C.c_ulonglong, C.c_void_p, C.c_longlong, C.c_void_p],
})

_TYPESTR = {torch.float32: "<f4", torch.int32: "<i4", torch.int64: "<i8", torch.uint8: "|u1", torch.int16: "<i2"}


class _RawCuda:
    def __init__(self, ptr, numel, typestr):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 2}


def tensor_from_ptr(ptr, numel, dtype, device):
    """Wrap raw device memory as a torch tensor (no ownership)."""
    if dtype == torch.bfloat16:
        return torch.as_tensor(_RawCuda(ptr, numel, "<i2"), device=device).view(torch.bfloat16)
    return torch.as_tensor(_RawCuda(ptr, numel, _TYPESTR[dtype]), device=device)


def ptr_array(ptrs):
    return (C.c_void_p * len(ptrs))(*[p if p else None for p in ptrs])


class SymmBuffer:
    """One slab per rank; `ptrs[r]` is rank r's slab as seen from this process."""

    def __init__(self, ptrs, nbytes, rank, device, owner, kind="ipc", mapped=None):
        self.ptrs, self.nbytes, self.rank, self.device, self._owner = list(ptrs), nbytes, rank, device, owner
        self.kind, self.mapped = kind, mapped or [nbytes] * len(self.ptrs)
        self.mc_ptr = None      # NVLS multicast address of the same offsets on every rank (SymmetricHeap.alloc(multicast=True))

    @property
    def local_ptr(self):
        return self.ptrs[self.rank]

    def tensor(self, dtype, numel=None, offset_bytes=0):
        esize = torch.empty(0, dtype=dtype).element_size()
        numel = (self.nbytes - offset_bytes) // esize if numel is None else numel
        return tensor_from_ptr(self.local_ptr + offset_bytes, numel, dtype, self.device)

    def peer_ptrs(self, offset_bytes=0):
        return [p + offset_bytes for p in self.ptrs]


class _FdExchange:
    """Pass file descriptors between the ranks of one box (SCM_RIGHTS over abstract unix sockets)."""

    def __init__(self, ctx):
        import os
        import socket

        self.ctx, self.socket = ctx, socket
        self.tag = f"zoo-symm-{os.environ.get('MASTER_PORT', '0')}-{os.getppid() if os.environ.get('TORCHELASTIC_RUN_ID') else 0}"
        self.listener = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self.listener.bind("\0" + self._name(ctx.rank))
        self.listener.listen(64)

    def _name(self, rank):
        return f"{self.tag}-{rank}"

    def exchange(self, fd, payload: bytes):
        """Send (fd, payload) to every peer; returns {rank: (fd, payload)} received from the peers."""
        import torch.distributed as dist

        W, r, socket = self.ctx.world_size, self.ctx.rank, self.socket
        dist.barrier()            # every listener exists
        for p in range(W):
            if p == r:
                continue
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect("\0" + self._name(p))
            socket.send_fds(c, [r.to_bytes(4, "little") + payload], [fd])
            c.close()
        got = {}
        while len(got) < W - 1:
            conn, _ = self.listener.accept()
            msg, fds, _, _ = socket.recv_fds(conn, 4096, 1)
            conn.close()
            got[int.from_bytes(msg[:4], "little")] = (fds[0], msg[4:])
        dist.barrier()
        return got

    def close(self):
        self.listener.close()


class SymmetricHeap:
    """`alloc` is collective: every rank calls it with the same size, in the same order.

    Two transports: ``vmm`` (cuMemCreate + file-descriptor export, mapped by peers with large pages -- the
    default; needed for random access into multi-GB peer buffers, see csrc/comm/symm_vmm.cu) and ``ipc``
    (cudaMalloc + cudaIpc handles; fallback, select with ZOO_SYMM=ipc)."""

    def __init__(self, ctx):
        import os

        self.ctx = ctx
        self.buffers = []
        self.lib = _nl.cuda_lib()
        _nl._bind(self.lib, _nl._CUDA_SIGS)
        self.mode = os.environ.get("ZOO_SYMM", "vmm")
        self._fdx = None

    def multicast_granularity(self):
        """Size granularity of NVLS multicast objects for this team, or 0 when some rank cannot use them (decided collectively,
        once).  ZOO_NVLS=0 switches multicast off."""
        import os

        if getattr(self, "_mc_gran", None) is None:
            g = 0
            if self.ctx.world_size > 1 and self.mode == "vmm" and os.environ.get("ZOO_NVLS", "auto") != "0":
                import torch.distributed as dist

                ok, gran = C.c_int(0), C.c_longlong(0)
                try:
                    self.lib.zoo_mc_supported(C.byref(ok))
                    if ok.value and self.lib.zoo_mc_granularity(self.ctx.world_size, C.byref(gran)) != 0:
                        ok.value = 0
                except Exception:
                    ok.value = 0
                t = torch.tensor([gran.value if ok.value else 0], dtype=torch.int64, device=self.ctx.device)
                dist.all_reduce(t, op=dist.ReduceOp.MIN)
                g = int(t)
            self._mc_gran = g
        return self._mc_gran

    def _agree(self, ok):
        import torch.distributed as dist

        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.ctx.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(int(t))

    def _bind_multicast(self, buf):
        """Create one multicast object over `buf` on every rank (collective); on any failure the buffer stays an ordinary symmetric
        buffer (mc_ptr None) on ALL ranks."""
        import logging
        import os

        ctx, W = self.ctx, self.ctx.world_size
        mapped = buf.mapped[ctx.rank]
        handle, mfd, err = C.c_ulonglong(0), C.c_int(-1), None
        if ctx.rank == 0:
            if self.lib.zoo_mc_create(mapped, W, C.byref(handle), C.byref(mfd)) != 0:
                err = self.lib.zoo_last_error().decode(errors="replace")
        if not self._agree(err is None):
            logging.getLogger("zoo").warning("NVLS multicast object could not be created (%s); using peer loads / stores", err)
            return
        send_fd = mfd.value if ctx.rank == 0 else os.open(os.devnull, os.O_RDONLY)
        got = self._fdx.exchange(send_fd, b"mc")
        if ctx.rank != 0:
            os.close(send_fd)
            if self.lib.zoo_mc_import(got[0][0], C.byref(handle)) != 0:
                err = self.lib.zoo_last_error().decode(errors="replace")
        for r, (pfd, _) in got.items():
            self.lib.zoo_close_fd(pfd)
        if ctx.rank == 0:
            self.lib.zoo_close_fd(mfd.value)
        if err is None and self.lib.zoo_mc_add_device(handle) != 0:
            err = self.lib.zoo_last_error().decode(errors="replace")
        if not self._agree(err is None):        # (also the barrier: every device is added before anyone binds)
            logging.getLogger("zoo").warning("NVLS multicast set-up failed (%s); using peer loads / stores", err)
            return
        mc = C.c_void_p()
        if self.lib.zoo_mc_bind(handle, C.c_void_p(buf.local_ptr), mapped, C.byref(mc)) != 0:
            err = self.lib.zoo_last_error().decode(errors="replace")
        if not self._agree(err is None):
            logging.getLogger("zoo").warning("NVLS multicast bind failed (%s); using peer loads / stores", err)
            return
        buf.mc_ptr = mc.value

    def _alloc_vmm(self, nbytes, align=0):
        ctx = self.ctx
        ptr, fd, mapped = C.c_void_p(), C.c_int(-1), C.c_longlong(0)
        if align:
            _nl.check(self.lib.zoo_vmm_alloc_aligned(nbytes, align, C.byref(ptr), C.byref(fd), C.byref(mapped)), "zoo_vmm_alloc_aligned")
        else:
            _nl.check(self.lib.zoo_vmm_alloc(nbytes, C.byref(ptr), C.byref(fd), C.byref(mapped)), "zoo_vmm_alloc")
        if ctx.world_size == 1:
            self.lib.zoo_close_fd(fd.value)
            return SymmBuffer([ptr.value], nbytes, 0, ctx.device, self, "vmm", [mapped.value])
        if self._fdx is None:
            self._fdx = _FdExchange(ctx)
        got = self._fdx.exchange(fd.value, mapped.value.to_bytes(8, "little"))
        self.lib.zoo_close_fd(fd.value)
        ptrs, sizes = [None] * ctx.world_size, [0] * ctx.world_size
        ptrs[ctx.rank], sizes[ctx.rank] = ptr.value, mapped.value
        for r, (pfd, payload) in got.items():
            p, sz = C.c_void_p(), int.from_bytes(payload[:8], "little")
            _nl.check(self.lib.zoo_vmm_import(pfd, sz, C.byref(p)), "zoo_vmm_import")
            self.lib.zoo_close_fd(pfd)
            ptrs[r], sizes[r] = p.value, sz
        return SymmBuffer(ptrs, nbytes, ctx.rank, ctx.device, self, "vmm", sizes)

    def alloc(self, nbytes, multicast=False) -> SymmBuffer:
        """`multicast=True` additionally binds the buffer to an NVLS multicast object when every rank supports it
        (`buf.mc_ptr`; None otherwise)."""
        import torch.distributed as dist

        nbytes = (int(nbytes) + 255) // 256 * 256
        if self.mode == "vmm":
            try:
                gran = self.multicast_granularity() if multicast else 0
                buf = self._alloc_vmm(nbytes, gran)
                self.buffers.append(buf)
                if gran:
                    self._bind_multicast(buf)
                return buf
            except _nl.NativeError as e:
                if self.buffers or self.ctx.world_size > 1:
                    raise
                import logging

                logging.getLogger("zoo").warning("VMM symmetric memory unavailable (%s); using CUDA IPC", e)
                self.mode = "ipc"
        ptr = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        _nl.check(self.lib.zoo_symm_alloc(nbytes, C.byref(ptr), handle), "zoo_symm_alloc")
        ctx = self.ctx
        if ctx.world_size == 1:
            buf = SymmBuffer([ptr.value], nbytes, 0, ctx.device, self)
            self.buffers.append(buf)
            return buf
        handles = [None] * ctx.world_size
        dist.all_gather_object(handles, bytes(handle))
        ptrs = []
        for r, h in enumerate(handles):
            if r == ctx.rank:
                ptrs.append(ptr.value)
                continue
            p = C.c_void_p()
            hb = (C.c_ubyte * 64).from_buffer_copy(h)
            _nl.check(self.lib.zoo_symm_open(hb, C.byref(p)), "zoo_symm_open")
            ptrs.append(p.value)
        buf = SymmBuffer(ptrs, nbytes, ctx.rank, ctx.device, self)
        self.buffers.append(buf)
        return buf

    def close(self):
        torch.cuda.synchronize()
        if self.ctx.world_size > 1:
            import torch.distributed as dist

            if dist.is_initialized():
                dist.barrier()
        for b in self.buffers:
            for r, p in enumerate(b.ptrs):
                if not p:
                    continue
                if b.kind == "vmm":
                    self.lib.zoo_vmm_free(C.c_void_p(p), b.mapped[r])
                elif r != b.rank:
                    self.lib.zoo_symm_close(C.c_void_p(p))
        if self.ctx.world_size > 1:
            import torch.distributed as dist

            if dist.is_initialized():
                dist.barrier()
        for b in self.buffers:
            if b.kind != "vmm":
                self.lib.zoo_symm_free(C.c_void_p(b.local_ptr))
        self.buffers = []
        if self._fdx is not None:
            self._fdx.close()
            self._fdx = None


_HEAP = None


def get_heap(ctx) -> SymmetricHeap:
    global _HEAP
    if _HEAP is None or _HEAP.ctx is not ctx:
        _HEAP = SymmetricHeap(ctx)
    return _HEAP
