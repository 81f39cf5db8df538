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

    def __init__(self, ptrs, nbytes, rank, device, owner):
        self.ptrs, self.nbytes, self.rank, self.device, self._owner = list(ptrs), nbytes, rank, device, owner

    @property
    def local_ptr(self):
        return self.ptrs[self.rank]

    def tensor(self, dtype, numel=None, offset_bytes=0):
        esize = torch.empty(0, dtype=dtype).element_size()
        numel = (self.nbytes - offset_bytes) // esize if numel is None else numel
        return tensor_from_ptr(self.local_ptr + offset_bytes, numel, dtype, self.device)

    def peer_ptrs(self, offset_bytes=0):
        return [p + offset_bytes for p in self.ptrs]


class SymmetricHeap:
    def __init__(self, ctx):
        self.ctx = ctx
        self.buffers = []
        self.lib = _nl.cuda_lib()
        _nl._bind(self.lib, _nl._CUDA_SIGS)

    def alloc(self, nbytes) -> SymmBuffer:
        import torch.distributed as dist

        nbytes = (int(nbytes) + 255) // 256 * 256
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
        for b in self.buffers:
            for r, p in enumerate(b.ptrs):
                if r != b.rank and p:
                    self.lib.zoo_symm_close(C.c_void_p(p))
        if self.ctx.world_size > 1:
            import torch.distributed as dist

            if dist.is_initialized():
                dist.barrier()
        for b in self.buffers:
            self.lib.zoo_symm_free(C.c_void_p(b.local_ptr))
        self.buffers = []


_HEAP = None


def get_heap(ctx) -> SymmetricHeap:
    global _HEAP
    if _HEAP is None or _HEAP.ctx is not ctx:
        _HEAP = SymmetricHeap(ctx)
    return _HEAP
