"""Off-heap arrays for cached training data.

Reference: zoo/.../feature/pmem/{NativeArray,VarLenBytesArray,FixLenBytesArray,FloatArray,VarLenFloatsArray}.scala and
`MemoryType` (feature/pmem/MemoryType.scala: DRAM / PMEM / DIRECT / DISK_AND_DRAM(n)); the allocator is memkind over Optane
persistent memory through JNI (PersistentMemoryAllocator.java:37-43), configured by `analytics.zoo.pmem.paths`.  There is no
Optane tier on a B200 box.  The arrays keep their interface over the C++ arena (`csrc/runtime/arena.cpp`): anonymous host
memory for DIRECT, a file-backed mapping when a path is given for PMEM (an NVMe / tmpfs path plays the role of the pmem mount),
optionally page-locked so that batches assembled from it can be copied to the GPU asynchronously.
"""
from __future__ import annotations

import numpy as np

from ..native.rt import Arena, RecordStore


class MemoryType:
    DRAM, PMEM, DIRECT = "DRAM", "PMEM", "DIRECT"

    @staticmethod
    def DISK_AND_DRAM(num_slice):
        return f"DISK_AND_DRAM({int(num_slice)})"


def _arena(total_bytes, memory_type, path):
    if memory_type not in (MemoryType.PMEM, MemoryType.DIRECT, MemoryType.DRAM):
        raise ValueError(f"unsupported memory type {memory_type!r} for a native array")
    if memory_type == MemoryType.PMEM and path is not None:
        return Arena(int(total_bytes), kind="file", path=path)
    return Arena(int(total_bytes), kind="host")


class NativeArray:
    """A block of `total_bytes` off-heap bytes; `free()` releases it (NativeArray.scala:38-100)."""

    def __init__(self, total_bytes, memory_type=MemoryType.DIRECT, path=None):
        self.total_bytes, self.memory_type = int(total_bytes), memory_type
        self.arena = _arena(max(self.total_bytes, 64) + 4096, memory_type, path)
        self.deleted = False

    def pin(self):
        """Page-lock the block (cudaHostRegister) so H2D copies from it are asynchronous; a no-op without CUDA."""
        try:
            self.arena.pin()
        except Exception:  # noqa: BLE001 - pinning is an optimisation
            pass
        return self

    def free(self):
        if not self.deleted:
            self.arena.close()
            self.deleted = True

    def _check(self):
        if self.deleted:
            raise RuntimeError("the native array has been freed")


class FixLenBytesArray(NativeArray):
    """`record_num` records of `record_bytes` bytes each."""

    def __init__(self, record_num, record_bytes, memory_type=MemoryType.DIRECT, path=None):
        super().__init__(int(record_num) * int(record_bytes), memory_type, path)
        self.record_num, self.record_bytes = int(record_num), int(record_bytes)
        self.data = self.arena.ndarray((self.record_num, self.record_bytes), np.uint8)

    def set(self, i, value):
        self._check()
        b = np.frombuffer(bytes(value), np.uint8)
        if b.size != self.record_bytes:
            raise ValueError(f"record {i}: expected {self.record_bytes} bytes, got {b.size}")
        self.data[i] = b

    def get(self, i):
        self._check()
        return self.data[i].tobytes()


class FloatArray(NativeArray):
    """`record_num` float32 values (`FloatArray(recordNum)`); `FixLenFloatsArray`-style rows with `record_len`."""

    def __init__(self, record_num, record_len=1, memory_type=MemoryType.DIRECT, path=None):
        super().__init__(int(record_num) * int(record_len) * 4, memory_type, path)
        self.record_num, self.record_len = int(record_num), int(record_len)
        self.data = self.arena.ndarray((self.record_num, self.record_len), np.float32)

    def set(self, i, value):
        self._check()
        self.data[i] = np.asarray(value, np.float32).reshape(self.record_len)

    def get(self, i):
        self._check()
        return self.data[i, 0] if self.record_len == 1 else self.data[i].copy()

    def batch(self, idx):
        """Rows `idx` as one contiguous float32 array (a training batch)."""
        self._check()
        return self.data[np.asarray(idx, dtype=np.int64)]


class VarLenBytesArray(NativeArray):
    """Variable-length records appended in order and read back by index (JPEG bytes of an ImageNet cache)."""

    def __init__(self, record_num, total_bytes, memory_type=MemoryType.DIRECT, path=None):
        super().__init__(int(total_bytes) + 16 * int(record_num), memory_type, path)
        self.record_num = int(record_num)
        self.store = RecordStore(self.arena)

    def set(self, i, value):
        self._check()
        if i != len(self.store):
            raise ValueError(f"records are appended in order: expected index {len(self.store)}, got {i}")
        self.store.append(bytes(value))

    append = lambda self, value: self.set(len(self.store), value)   # noqa: E731

    def get(self, i):
        self._check()
        return self.store.get(i)

    def __len__(self):
        return len(self.store)


class VarLenFloatsArray(VarLenBytesArray):
    """Variable-length float32 records."""

    def __init__(self, record_num, total_floats, memory_type=MemoryType.DIRECT, path=None):
        super().__init__(record_num, int(total_floats) * 4, memory_type, path)

    def set(self, i, value):
        super().set(i, np.asarray(value, np.float32).tobytes())

    def get(self, i):
        return np.frombuffer(super().get(i), np.float32)
