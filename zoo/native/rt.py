"""Python faces of the C++ runtime (``libzoo_rt.so``): stream broker, arena allocator, record
store, CRC32C / TFRecord framing.  See csrc/runtime/*.cpp for what each replaces in the reference."""
from __future__ import annotations

import ctypes as C
import socket
import threading

import numpy as np

from . import lib as _nl

_V, _I, _L = C.c_void_p, C.c_int, C.c_longlong
_bound = False


def rt():
    global _bound
    lib = _nl.rt_lib()
    if not _bound:
        sig = {
            "zoo_broker_create": ([_L], _V), "zoo_broker_destroy": ([_V], None), "zoo_broker_serve": ([_V, C.c_char_p, _I], _I),
            "zoo_broker_stop_server": ([_V], None), "zoo_broker_used_memory": ([_V], _L),
            "zoo_broker_command": ([_V, _I, C.POINTER(C.c_char_p), C.POINTER(_L), C.POINTER(C.POINTER(C.c_char)), C.POINTER(_L)], _I),
            "zoo_broker_free": ([C.POINTER(C.c_char)], None),
            "zoo_broker_read_pack": ([_V, C.c_char_p, C.c_char_p, _L, _L, C.c_char_p, _V, _L, C.POINTER(_L), _V, _L, C.POINTER(_L)], _L),
            "zoo_broker_hset_topn": ([_V, C.c_char_p, C.c_char_p, _L, _V, _L, _I, _V, _V], _L),
            "zoo_broker_hset_many": ([_V, C.c_char_p, C.c_char_p, _L, C.c_char_p, _V, C.c_char_p, _V], _L),
            "zoo_broker_xadd_many": ([_V, C.c_char_p, _L, C.c_char_p, _V, C.c_char_p, _V, _L, _I, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p)], _L),
            "zoo_arena_create": ([_I, _L, C.c_char_p], _V), "zoo_arena_destroy": ([_V], None), "zoo_arena_base": ([_V], _V),
            "zoo_arena_capacity": ([_V], _L), "zoo_arena_used": ([_V], _L), "zoo_arena_reset": ([_V], None),
            "zoo_arena_alloc": ([_V, _L, _I], _L), "zoo_arena_evict": ([_V, _L, _L], _I),
            "zoo_records_create": ([_V], _V), "zoo_records_destroy": ([_V], None), "zoo_records_count": ([_V], _L),
            "zoo_records_append": ([_V, _V, _L], _L), "zoo_records_get": ([_V, _L, C.POINTER(_V)], _L),
            "zoo_records_gather": ([_V, _V, _L, _V, _L, _V, _I], _L),
            "zoo_b64_decode_batch": ([_V, _V, _I, _V, _V, _V, _I], _I),
            "zoo_crc32c": ([_V, _L], C.c_uint), "zoo_masked_crc32c": ([_V, _L], C.c_uint), "zoo_tfrecord_frame": ([_V, _L, _V], _L),
        }
        for name, (args, res) in sig.items():
            fn = getattr(lib, name)
            fn.argtypes, fn.restype = args, res
        _bound = True
    return lib


# ------------------------------------------------------------------------------- RESP --
class RespError(Exception):
    pass


def _to_bytes(a):
    if isinstance(a, bytes):
        return a
    if isinstance(a, (bytearray, memoryview)):
        return bytes(a)
    return str(a).encode()


def encode_command(args):
    out = [b"*%d\r\n" % len(args)]
    for a in args:
        b = _to_bytes(a)
        out.append(b"$%d\r\n" % len(b))
        out.append(b)
        out.append(b"\r\n")
    return b"".join(out)


def parse_reply(buf: bytes, pos: int = 0):
    """Parse one RESP2 value at `pos`; returns (value, next_pos) or raises IndexError when incomplete."""
    t = buf[pos:pos + 1]
    end = buf.index(b"\r\n", pos)
    line = buf[pos + 1:end]
    nxt = end + 2
    if t == b"+":
        return line.decode(), nxt
    if t == b"-":
        return RespError(line.decode()), nxt
    if t == b":":
        return int(line), nxt
    if t == b"$":
        n = int(line)
        if n < 0:
            return None, nxt
        if nxt + n + 2 > len(buf):
            raise IndexError("incomplete")
        return buf[nxt:nxt + n], nxt + n + 2
    if t == b"*":
        n = int(line)
        if n < 0:
            return None, nxt
        items = []
        for _ in range(n):
            v, nxt = parse_reply(buf, nxt)
            items.append(v)
        return items, nxt
    raise RespError(f"bad reply type {t!r}")


class _Commands:
    """Redis-style convenience verbs on top of `execute(*args)` (the subset the reference's client and
    server use: pyzoo/zoo/serving/client.py:87-151, ClusterServing.scala:105-140,276-310)."""

    def execute(self, *args):  # pragma: no cover - interface
        raise NotImplementedError

    def ping(self):
        return self.execute("PING") == "PONG"

    def xadd(self, name, fields: dict, id="*", maxlen=None):
        args = ["XADD", name] + (["MAXLEN", maxlen] if maxlen is not None else []) + [id]
        for k, v in fields.items():
            args += [k, v]
        return self.execute(*args)

    def xlen(self, name):
        return self.execute("XLEN", name)

    def xtrim(self, name, maxlen, approximate=False):
        return self.execute("XTRIM", name, "MAXLEN", maxlen)

    def xgroup_create(self, name, group, id="$", mkstream=False):
        return self.execute("XGROUP", "CREATE", name, group, id, *(["MKSTREAM"] if mkstream else []))

    def xreadgroup(self, group, consumer, streams: dict, count=None, block=None, noack=False):
        args = ["XREADGROUP", "GROUP", group, consumer]
        if count is not None:
            args += ["COUNT", count]
        if block is not None:
            args += ["BLOCK", block]
        if noack:
            args += ["NOACK"]
        args += ["STREAMS"] + list(streams.keys()) + list(streams.values())
        return self._streams(self.execute(*args))

    def xread(self, streams: dict, count=None, block=None):
        args = ["XREAD"]
        if count is not None:
            args += ["COUNT", count]
        if block is not None:
            args += ["BLOCK", block]
        args += ["STREAMS"] + list(streams.keys()) + list(streams.values())
        return self._streams(self.execute(*args))

    @staticmethod
    def _streams(r):
        if r is None:
            return []
        out = []
        for name, entries in r:
            out.append((name.decode(), [(eid.decode(), {f[i].decode(): f[i + 1] for i in range(0, len(f), 2)}) for eid, f in entries]))
        return out

    def xack(self, name, group, *ids):
        return self.execute("XACK", name, group, *ids)

    def hset(self, name, key=None, value=None, mapping=None):
        args = ["HSET", name]
        if key is not None:
            args += [key, value]
        for k, v in (mapping or {}).items():
            args += [k, v]
        return self.execute(*args)

    def hgetall(self, name):
        r = self.execute("HGETALL", name)
        return {r[i]: r[i + 1] for i in range(0, len(r), 2)}

    def keys(self, pattern="*"):
        return self.execute("KEYS", pattern)

    def delete(self, *names):
        return self.execute("DEL", *names)

    def info(self, section=None):
        raw = self.execute("INFO", *([section] if section else []))
        out = {}
        for line in raw.decode().splitlines():
            if ":" in line and not line.startswith("#"):
                k, v = line.split(":", 1)
                out[k] = int(v) if v.lstrip("-").isdigit() else v
        return out

    def config_set(self, name, value):
        return self.execute("CONFIG", "SET", name, value)

    def flushall(self):
        return self.execute("FLUSHALL")


class Broker(_Commands):
    """In-process stream broker; `serve()` additionally exposes it on a RESP socket."""

    def __init__(self, maxmemory=0):
        self._lib = rt()
        self._h = self._lib.zoo_broker_create(int(maxmemory))
        self.port = None

    def serve(self, host="127.0.0.1", port=0):
        p = self._lib.zoo_broker_serve(self._h, host.encode(), int(port))
        if p < 0:
            raise OSError(f"broker could not listen on {host}:{port}")
        self.port = p
        return p

    def stop_server(self):
        if self._h:
            self._lib.zoo_broker_stop_server(self._h)
            self.port = None

    def execute(self, *args):
        bs = [_to_bytes(a) for a in args]
        n = len(bs)
        argv = (C.c_char_p * n)(*bs)
        lens = (_L * n)(*[len(b) for b in bs])
        out, outlen = C.POINTER(C.c_char)(), _L()
        rc = self._lib.zoo_broker_command(self._h, n, argv, lens, C.byref(out), C.byref(outlen))
        if rc != 0:
            raise MemoryError("broker reply allocation failed")
        try:
            reply = C.string_at(out, outlen.value)
        finally:
            self._lib.zoo_broker_free(out)
        v, _ = parse_reply(reply)
        if isinstance(v, RespError):
            raise v
        return v

    def used_memory(self):
        return self._lib.zoo_broker_used_memory(self._h)

    # ---- batch paths (one native call per batch; used by the in-process serving engine and bulk producers) --------
    def read_pack(self, stream, group, count, field, block_ms=5, buf=None, uri_buf=None, raw_uris=False):
        """Up to `count` new entries whose `field` values have equal length -> (n, uint8 array [n, rec_len], list of uris)."""
        if buf is None:
            buf = np.empty(64 << 20, dtype=np.uint8)
        if uri_buf is None:
            uri_buf = np.empty(max(1 << 20, count * 64), dtype=np.uint8)
        rec_len, uris_len = _L(0), _L(0)
        n = self._lib.zoo_broker_read_pack(self._h, stream.encode(), group.encode(), int(count), int(block_ms), field.encode(), buf.ctypes.data,
                                           buf.nbytes, C.byref(rec_len), uri_buf.ctypes.data, uri_buf.nbytes, C.byref(uris_len))
        if n < 0:
            raise RespError(f"NOGROUP No such consumer group '{group}' for key '{stream}'")
        if n == 0:
            return 0, None, []
        blob = uri_buf[:uris_len.value].tobytes()
        uris = blob if raw_uris else blob.decode().split("\n")[:-1]
        return n, buf[:n * rec_len.value].reshape(n, rec_len.value), uris

    def hset_topn(self, prefix, field, uris_blob: bytes, idx: np.ndarray, val: np.ndarray):
        """Top-N results of a batch: `uris_blob` is the newline-separated uri blob of `read_pack(raw_uris=True)`, idx / val are
        [n, k]; the "[[idx,prob],...]" strings are formatted natively."""
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        val = np.ascontiguousarray(val, dtype=np.float32)
        n, k = idx.shape
        r = self._lib.zoo_broker_hset_topn(self._h, prefix.encode(), field.encode(), n, uris_blob, len(uris_blob), k, idx.ctypes.data, val.ctypes.data)
        if r < 0:
            raise RespError("OOM command not allowed when used memory > 'maxmemory'.")
        return r

    def hset_many(self, prefix, field, keys, values):
        kb = [k.encode() if isinstance(k, str) else k for k in keys]
        vb = [v.encode() if isinstance(v, str) else v for v in values]
        koff = np.zeros(len(kb) + 1, dtype=np.int64); np.cumsum([len(k) for k in kb], out=koff[1:])
        voff = np.zeros(len(vb) + 1, dtype=np.int64); np.cumsum([len(v) for v in vb], out=voff[1:])
        n = self._lib.zoo_broker_hset_many(self._h, prefix.encode(), field.encode(), len(kb), b"".join(kb), koff.ctypes.data, b"".join(vb), voff.ctypes.data)
        if n < 0:
            raise RespError("OOM command not allowed when used memory > 'maxmemory'.")
        return n

    def xadd_many(self, stream, uris, field, records: np.ndarray, extra=None):
        """One entry per row of `records` ([n, ...] array; raw little-endian bytes), fields {uri, field, **extra}."""
        recs = np.ascontiguousarray(records)
        n = recs.shape[0]
        ub = [u.encode() for u in uris]
        uoff = np.zeros(n + 1, dtype=np.int64); np.cumsum([len(u) for u in ub], out=uoff[1:])
        extra = extra or {}
        names = (C.c_char_p * len(extra))(*[k.encode() for k in extra])
        vals = (C.c_char_p * len(extra))(*[str(v).encode() for v in extra.values()])
        return self._lib.zoo_broker_xadd_many(self._h, stream.encode(), n, b"".join(ub), uoff.ctypes.data, field.encode(), recs.ctypes.data,
                                              recs.nbytes // max(n, 1), len(extra), names, vals)

    def close(self):
        if self._h:
            self._lib.zoo_broker_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RespClient(_Commands):
    """Minimal RESP2 socket client (works against the broker's listener or a real Redis)."""

    def __init__(self, host="127.0.0.1", port=6379, timeout=None):
        self.sock = socket.create_connection((host, int(port)), timeout=timeout)
        self.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self._buf = b""
        self._lock = threading.Lock()      # one request / reply at a time: a serving engine calls from several threads

    def execute(self, *args):
        with self._lock:
            self.sock.sendall(encode_command(args))
            while True:
                try:
                    v, nxt = parse_reply(self._buf)
                    self._buf = self._buf[nxt:]
                    if isinstance(v, RespError):
                        raise v
                    return v
                except (ValueError, IndexError):
                    chunk = self.sock.recv(1 << 20)
                    if not chunk:
                        raise ConnectionError("broker closed the connection")
                    self._buf += chunk

    def close(self):
        try:
            self.sock.close()
        except OSError:
            pass


# ------------------------------------------------------------------- arena / records --
class Arena:
    """Page-aligned bump arena: kind 'host' (anonymous memory) or 'file' (mmap of `path`)."""

    def __init__(self, capacity, kind="host", path=None):
        self._lib = rt()
        self._h = self._lib.zoo_arena_create(1 if kind == "file" else 0, int(capacity), path.encode() if path else None)
        if not self._h:
            raise MemoryError(f"arena of {capacity} bytes ({kind}) could not be created")
        self.kind, self.capacity, self._pinned = kind, int(capacity), False

    @property
    def base(self):
        return self._lib.zoo_arena_base(self._h)

    @property
    def used(self):
        return self._lib.zoo_arena_used(self._h)

    def alloc(self, nbytes, align=64):
        off = self._lib.zoo_arena_alloc(self._h, int(nbytes), int(align))
        if off < 0:
            raise MemoryError(f"arena full: {self.used} + {nbytes} > {self.capacity}")
        return off

    def view(self, offset, nbytes):
        """numpy uint8 view (no copy)."""
        return np.ctypeslib.as_array((C.c_ubyte * int(nbytes)).from_address(self.base + int(offset)))

    def ndarray(self, shape, dtype):
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        off = self.alloc(n, 256)
        return self.view(off, n).view(dtype).reshape(shape)

    def pin(self):
        """Page-lock the arena for async DMA (cudaHostRegister through torch's cudart)."""
        import torch

        if self._pinned or not torch.cuda.is_available():
            return self._pinned
        rc = torch.cuda.cudart().cudaHostRegister(self.base, self.capacity, 0)
        self._pinned = int(rc) == 0
        return self._pinned

    def evict(self, offset, nbytes):
        return self._lib.zoo_arena_evict(self._h, int(offset), int(nbytes))

    def reset(self):
        self._lib.zoo_arena_reset(self._h)

    def close(self):
        if self._h:
            if self._pinned:
                import torch

                torch.cuda.cudart().cudaHostUnregister(self.base)
            self._lib.zoo_arena_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RecordStore:
    """Variable-length byte records in an arena (the reference's VarLenBytesArray)."""

    def __init__(self, arena: Arena):
        self.arena = arena
        self._lib = rt()
        self._h = self._lib.zoo_records_create(arena._h)

    def append(self, data) -> int:
        b = _to_bytes(data) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data).tobytes()
        i = self._lib.zoo_records_append(self._h, b, len(b))
        if i < 0:
            raise MemoryError("record store arena is full")
        return i

    def __len__(self):
        return self._lib.zoo_records_count(self._h)

    def get(self, i) -> bytes:
        p = _V()
        n = self._lib.zoo_records_get(self._h, int(i), C.byref(p))
        if n < 0:
            raise IndexError(i)
        return C.string_at(p.value, n)

    def gather(self, idx, out: np.ndarray | None = None, threads=4):
        """Copy records `idx` back to back -> (uint8 buffer, offsets[n+1])."""
        idx = np.ascontiguousarray(idx, dtype=np.int64)
        offs = np.empty(len(idx) + 1, dtype=np.int64)
        if out is None:
            p = _V()
            total = sum(self._lib.zoo_records_get(self._h, int(i), C.byref(p)) for i in idx)
            out = np.empty(max(total, 1), dtype=np.uint8)
        n = self._lib.zoo_records_gather(self._h, idx.ctypes.data, len(idx), out.ctypes.data, out.nbytes, offs.ctypes.data, threads)
        if n < 0:
            raise ValueError("gather: destination too small or index out of range")
        return out[:n], offs

    def close(self):
        if self._h:
            self._lib.zoo_records_destroy(self._h)
            self._h = None


def crc32c(data: bytes) -> int:
    return rt().zoo_crc32c(data, len(data))


def masked_crc32c(data: bytes) -> int:
    return rt().zoo_masked_crc32c(data, len(data))


def tfrecord_frame(data: bytes) -> bytes:
    out = C.create_string_buffer(len(data) + 16)
    n = rt().zoo_tfrecord_frame(data, len(data), out)
    return out.raw[:n]


def b64_decode_batch(strings, out: "np.ndarray | None" = None, threads=8):
    """Decode a list of base64 byte strings on native threads (no interpreter lock).  Returns (buffer uint8, offsets, lengths):
    record i is buffer[offsets[i] : offsets[i] + lengths[i]]; a malformed record has length -1."""
    n = len(strings)
    lens = np.fromiter((len(s) for s in strings), dtype=np.int64, count=n)
    caps = (lens + 3) // 4 * 3
    offs = np.zeros(n, dtype=np.int64)
    if n:
        np.cumsum((caps[:-1] + 15) // 16 * 16, out=offs[1:])
    total = int(offs[-1] + caps[-1]) if n else 0
    if out is None or out.size < total:
        out = np.empty(max(total, 1), dtype=np.uint8)
    out_len = np.zeros(n, dtype=np.int64)
    ptrs = (C.c_char_p * n)(*strings)
    rt().zoo_b64_decode_batch(C.cast(ptrs, _V), lens.ctypes.data, n, out.ctypes.data, offs.ctypes.data, out_len.ctypes.data, int(threads))
    return out, offs, out_len
