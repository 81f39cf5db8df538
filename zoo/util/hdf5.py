"""HDF5 files without libhdf5 / h5py: a reader with enough of the format for Keras weight files and MATLAB v7.3 `.mat` files, and a
writer of the classic layout (`write_hdf5`).

Where the reference meets HDF5: `Net.load_keras(json_path, hdf5_path)` (Keras `save_weights` files, pyzoo/zoo/pipeline/api/net/
net_load.py:117-140) and the 3-D image tests (`image3d/a.mat`, a MATLAB v7.3 file, read there with h5py).  Covered: superblock
versions 0-1 with version-1 object headers (what h5py and MATLAB write by default), groups through symbol tables (v1 B-trees + local
heaps), datasets with compact, contiguous or chunked layout, the deflate / shuffle / fletcher32 filters, fixed-point, floating-point
and fixed-length string types, attributes (scalars, arrays, strings).  Not covered (raises): superblock 2/3 with version-2 object
headers and fractal-heap groups (`libver="latest"` files), variable-length types, compound types, external storage.

    with File(path) as f:
        f.keys(); f["group/dataset"][...]; f["dataset"].shape; f.attrs["layer_names"]
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class _Type:
    def __init__(self, dtype=None, string_len=None):
        self.dtype, self.string_len = dtype, string_len


class File:
    def __init__(self, path, mode="r"):
        if mode != "r":
            raise NotImplementedError("File() reads; write a whole file with zoo.util.hdf5.write_hdf5(path, tree, attrs)")
        with open(path, "rb") as f:
            self._b = f.read()
        base = next((o for o in (0, 512, 1024, 2048, 4096) if self._b[o:o + 8] == _SIG), None)
        if base is None:
            raise ValueError(f"{path} is not an HDF5 file")
        version = self._b[base + 8]
        if version > 1:
            raise NotImplementedError(f"HDF5 superblock version {version} (files written with libver='latest') is not read")
        if self._b[base + 13] != 8 or self._b[base + 14] != 8:
            raise NotImplementedError("HDF5 files with offsets / lengths that are not 8 bytes wide")
        p = base + 24 + (4 if version == 1 else 0)
        self._base = struct.unpack_from("<Q", self._b, p)[0]
        p += 32                                              # base, free-space info, end of file, driver info
        _, header, cache, _ = struct.unpack_from("<QQII", self._b, p)
        self.root = Group(self, header, "/")

    # ---- low level ---------------------------------------------------------------------------------------------------------------
    def _messages(self, addr):
        """[(type, flags, payload bytes)] of a version-1 object header, continuation blocks followed."""
        b = self._b
        addr += self._base
        if b[addr:addr + 4] == b"OHDR":
            raise NotImplementedError("version-2 object headers are not read")
        version, _, nmsg, _, size = struct.unpack_from("<BBHII", b, addr)
        if version != 1:
            raise NotImplementedError(f"object header version {version}")
        blocks, out = [(addr + 16, size)], []
        while blocks and len(out) < nmsg:
            pos, left = blocks.pop(0)
            end = pos + left
            while pos + 8 <= end and len(out) < nmsg:
                mtype, msize, flags = struct.unpack_from("<HHB", b, pos)
                body = b[pos + 8:pos + 8 + msize]
                pos += 8 + msize
                if mtype == 0x10:                              # continuation
                    off, ln = struct.unpack_from("<QQ", body, 0)
                    blocks.append((off + self._base, ln))
                out.append((mtype, flags, body))
        return out

    def _heap_string(self, heap_addr, offset):
        b = self._b
        h = heap_addr + self._base
        if b[h:h + 4] != b"HEAP":
            raise ValueError("bad local heap")
        data = struct.unpack_from("<Q", b, h + 24)[0] + self._base
        end = b.index(b"\x00", data + offset)
        return b[data + offset:end].decode()

    def _group_entries(self, btree, heap):
        """{name: object header address} of a symbol-table group."""
        b, out = self._b, {}

        def walk(addr):
            a = addr + self._base
            if b[a:a + 4] != b"TREE":
                raise ValueError("bad B-tree node")
            _, level, used = struct.unpack_from("<BBH", b, a + 4)
            p = a + 24
            children = []
            for _ in range(used):
                p += 8                                         # key
                children.append(struct.unpack_from("<Q", b, p)[0])
                p += 8
            for c in children:
                if level > 0:
                    walk(c)
                    continue
                s = c + self._base
                if b[s:s + 4] != b"SNOD":
                    raise ValueError("bad symbol node")
                n = struct.unpack_from("<H", b, s + 6)[0]
                for k in range(n):
                    name_off, header = struct.unpack_from("<QQ", b, s + 8 + 40 * k)
                    out[self._heap_string(heap, name_off)] = header

        walk(btree)
        return out

    # ---- message decoding --------------------------------------------------------------------------------------------------------
    @staticmethod
    def _dataspace(body):
        version, rank, flags = body[0], body[1], body[2]
        p = 8 if version == 1 else 4
        return tuple(struct.unpack_from("<Q", body, p + 8 * i)[0] for i in range(rank))

    @staticmethod
    def _datatype(body):
        cls, bits0 = body[0] & 0x0F, body[1]
        size = struct.unpack_from("<I", body, 4)[0]
        order = ">" if bits0 & 1 else "<"
        if cls == 0:
            return _Type(np.dtype(f"{order}{'i' if bits0 & 8 else 'u'}{size}"))
        if cls == 1:
            return _Type(np.dtype(f"{order}f{size}"))
        if cls == 3:
            return _Type(np.dtype(f"S{size}"), string_len=size)
        if cls == 4:                                           # bitfield
            return _Type(np.dtype(f"{order}u{size}"))
        if cls == 8:                                           # enum (MATLAB logicals): the base type follows
            return File._datatype(body[8:])
        if cls == 7:                                           # object reference: an address
            return _Type(np.dtype("<u8"))
        raise NotImplementedError(f"HDF5 datatype class {cls} (variable-length / compound / array types are not read)")

    @staticmethod
    def _filters(body):
        version, n = body[0], body[1]
        p = 8 if version == 1 else 2
        out = []
        for _ in range(n):
            fid = struct.unpack_from("<H", body, p)[0]
            if version == 1 or fid >= 256:
                name_len = struct.unpack_from("<H", body, p + 2)[0]
                _, nvals = struct.unpack_from("<HH", body, p + 4)
                p += 8 + ((name_len + 7) // 8 * 8 if version == 1 else name_len)
            else:
                _, nvals = struct.unpack_from("<HH", body, p + 2)
                p += 6
            vals = struct.unpack_from(f"<{nvals}I", body, p)
            p += 4 * nvals
            if version == 1 and nvals % 2:
                p += 4
            out.append((fid, vals))
        return out

    def _attribute(self, body):
        version = body[0]
        if version == 1:
            name_size, dt_size, ds_size = struct.unpack_from("<HHH", body, 2)
            pad = lambda n: (n + 7) // 8 * 8
            p = 8
            name = body[p:p + name_size].split(b"\x00")[0].decode()
            p += pad(name_size)
            dt, p = self._datatype(body[p:p + dt_size]), p + pad(dt_size)
            shape, p = self._dataspace(body[p:p + ds_size]), p + pad(ds_size)
        elif version in (2, 3):
            name_size, dt_size, ds_size = struct.unpack_from("<HHH", body, 2)
            p = 8 + (1 if version == 3 else 0)
            name = body[p:p + name_size].split(b"\x00")[0].decode()
            p += name_size
            dt, p = self._datatype(body[p:p + dt_size]), p + dt_size
            shape, p = self._dataspace(body[p:p + ds_size]), p + ds_size
        else:
            raise NotImplementedError(f"attribute message version {version}")
        n = int(np.prod(shape)) if shape else 1
        arr = np.frombuffer(body, dt.dtype, n, p).reshape(shape)
        return name, _decode(arr, dt)


def _decode(arr, dt):
    if dt.string_len is not None:
        strip = np.vectorize(lambda s: s.split(b"\x00")[0].rstrip(b" "), otypes=[object])
        out = strip(arr) if arr.shape else arr.reshape(1)[0].split(b"\x00")[0]
        return out if arr.shape else out
    out = arr.astype(arr.dtype.newbyteorder("="))
    return out if out.shape else out.reshape(1)[0]


class _Node:
    def __init__(self, file, header, name):
        self._f, self._header, self.name = file, header, name
        self._msgs = file._messages(header)

    @property
    def attrs(self):
        return dict(self._f._attribute(body) for t, _, body in self._msgs if t == 0x0C)


class Group(_Node):
    def __init__(self, file, header, name):
        super().__init__(file, header, name)
        sym = next((body for t, _, body in self._msgs if t == 0x11), None)
        self._entries = file._group_entries(*struct.unpack_from("<QQ", sym, 0)) if sym is not None else {}

    def keys(self):
        return list(self._entries)

    def __contains__(self, name):
        try:
            self[name]
            return True
        except KeyError:
            return False

    def __iter__(self):
        return iter(self._entries)

    def __len__(self):
        return len(self._entries)

    def __getitem__(self, path):
        node = self
        for part in [p for p in path.split("/") if p]:
            if not isinstance(node, Group) or part not in node._entries:
                raise KeyError(path)
            header = node._entries[part]
            msgs = node._f._messages(header)
            child = (node.name.rstrip("/") + "/" + part)
            node = Dataset(node._f, header, child) if any(t == 0x08 for t, _, _ in msgs) else Group(node._f, header, child)
        return node

    def items(self):
        return [(k, self[k]) for k in self._entries]

    def visit_datasets(self, prefix=""):
        """{path: Dataset} of everything below this group."""
        out = {}
        for k, v in self.items():
            if isinstance(v, Group):
                out.update(v.visit_datasets(prefix + k + "/"))
            else:
                out[prefix + k] = v
        return out


class Dataset(_Node):
    def __init__(self, file, header, name):
        super().__init__(file, header, name)
        get = lambda t: next((body for mt, _, body in self._msgs if mt == t), None)
        self.shape = File._dataspace(get(0x01))
        self._type = File._datatype(get(0x03))
        self._layout = get(0x08)
        self._pipeline = File._filters(get(0x0B)) if get(0x0B) is not None else []

    @property
    def dtype(self):
        return self._type.dtype.newbyteorder("=") if self._type.string_len is None else self._type.dtype

    def _raw(self):
        f, lay = self._f, self._layout
        b = f._b
        n = int(np.prod(self.shape)) if self.shape else 1
        item = self._type.dtype.itemsize
        version = lay[0]
        if version != 3:
            raise NotImplementedError(f"data layout message version {version}")
        cls = lay[1]
        if cls == 0:                                           # compact: the data sit in the header
            size = struct.unpack_from("<H", lay, 2)[0]
            return np.frombuffer(lay, self._type.dtype, n, 4).reshape(self.shape)
        if cls == 1:                                           # contiguous
            addr, size = struct.unpack_from("<QQ", lay, 2)
            if addr == _UNDEF:
                return np.zeros(self.shape, self._type.dtype)
            return np.frombuffer(b, self._type.dtype, n, addr + f._base).reshape(self.shape)
        if cls != 2:
            raise NotImplementedError(f"data layout class {cls}")
        rank1 = lay[2]                                         # dataset rank + 1 (the element size is the last "dimension")
        btree = struct.unpack_from("<Q", lay, 3)[0]
        chunk = struct.unpack_from(f"<{rank1}I", lay, 11)[:-1]
        out = np.zeros(self.shape, self._type.dtype)
        if btree == _UNDEF:
            return out

        def walk(addr):
            a = addr + f._base
            if b[a:a + 4] != b"TREE":
                raise ValueError("bad chunk B-tree node")
            _, level, used = struct.unpack_from("<BBH", b, a + 4)
            p = a + 24
            key = 8 + 8 * rank1
            for _ in range(used):
                csize, mask = struct.unpack_from("<II", b, p)
                offs = struct.unpack_from(f"<{rank1}Q", b, p + 8)[:-1]
                child = struct.unpack_from("<Q", b, p + key)[0]
                p += key + 8
                if level > 0:
                    walk(child)
                    continue
                raw = b[child + f._base:child + f._base + csize]
                for k, (fid, vals) in reversed(list(enumerate(self._pipeline))):
                    if mask & (1 << k):
                        continue                               # this filter was skipped for this chunk
                    if fid == 1:
                        raw = zlib.decompress(raw)
                    elif fid == 2:                             # shuffle: bytes of each element are stored plane by plane
                        cnt = len(raw) // item
                        raw = np.frombuffer(raw, np.uint8, cnt * item).reshape(item, cnt).T.tobytes() + raw[cnt * item:]
                    elif fid == 3:
                        raw = raw[:-4]                         # fletcher32 checksum at the end
                    else:
                        raise NotImplementedError(f"HDF5 filter {fid}")
                block = np.frombuffer(raw, self._type.dtype, int(np.prod(chunk))).reshape(chunk)
                sel = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, chunk, self.shape))
                out[sel] = block[tuple(slice(0, s.stop - s.start) for s in sel)]

        walk(btree)
        return out

    def __array__(self, dtype=None, copy=None):
        a = _decode(self._raw(), self._type)
        return np.asarray(a, dtype=dtype) if dtype is not None else np.asarray(a)

    def __getitem__(self, idx):
        return np.asarray(self)[idx]

    def __len__(self):
        return self.shape[0]


# File behaves like its root group
for _name in ("keys", "items", "__getitem__", "__contains__", "__iter__", "__len__", "visit_datasets"):
    setattr(File, _name, (lambda n: lambda self, *a, **k: getattr(self.root, n)(*a, **k))(_name))
File.attrs = property(lambda self: self.root.attrs)
File.__enter__ = lambda self: self
File.__exit__ = lambda self, *a: None
File.close = lambda self: None


# ---- writer (classic format: superblock 0, version-1 object headers, symbol-table groups, contiguous datasets) ---------------------
def _dt_message(dtype):
    dtype = np.dtype(dtype)
    if dtype.kind == "f":
        spec = {4: (0x1F, struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)), 8: (0x3F, struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023))}[dtype.itemsize]
        return struct.pack("<BBBBI", 0x11, 0x20, spec[0], 0, dtype.itemsize) + spec[1]
    if dtype.kind in "iu":
        return struct.pack("<BBBBI", 0x10, 0x08 if dtype.kind == "i" else 0, 0, 0, dtype.itemsize) + struct.pack("<HH", 0, 8 * dtype.itemsize)
    if dtype.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0, 0, 0, dtype.itemsize)
    raise TypeError(f"cannot store dtype {dtype}")


def _ds_message(shape):
    return struct.pack("<BBBBI", 1, len(shape), 0, 0, 0) + b"".join(struct.pack("<Q", int(s)) for s in shape)


def _pad8(b):
    return b + b"\x00" * (-len(b) % 8)


def _attr_message(name, value):
    if isinstance(value, str):
        value = value.encode()
    if isinstance(value, bytes):
        arr = np.array(value, dtype=f"S{max(len(value), 1)}")
    else:
        arr = np.asarray(value)
        if arr.dtype.kind == "U":
            arr = np.char.encode(arr, "utf-8")
        if arr.dtype.kind == "O":
            arr = np.array([v if isinstance(v, bytes) else str(v).encode() for v in arr.reshape(-1)]).reshape(arr.shape)
    nm = name.encode() + b"\x00"
    dt, ds = _dt_message(arr.dtype), _ds_message(arr.shape)
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + _pad8(nm) + _pad8(dt) + _pad8(ds) + np.ascontiguousarray(arr).tobytes()
    return 0x0C, body


def _object_header(messages):
    body = b""
    for mtype, payload in messages:
        payload = _pad8(payload)
        body += struct.pack("<HHBBBB", mtype, len(payload), 0, 0, 0, 0) + payload
    return struct.pack("<BBHII", 1, 0, len(messages), 1, len(body)) + b"\x00" * 4 + body


def write_hdf5(path, tree, attrs=None):
    """`tree`: {name: ndarray | nested dict}; attributes of a group under its key "__attrs__" ({name: value}), of the root in `attrs`.
    Writes the classic HDF5 layout (what h5py writes by default), contiguous datasets, no compression."""
    K_LEAF = 512                                             # symbols per leaf node = 2 * K_LEAF
    out = bytearray(b"\x00" * 96)                            # superblock placeholder (8-byte aligned start of data)

    def put(block):
        while len(out) % 8:
            out.append(0)
        addr = len(out)
        out.extend(block)
        return addr

    def write_dataset(arr):
        arr = np.ascontiguousarray(arr)
        if arr.dtype.kind == "U":
            arr = np.char.encode(arr, "utf-8")
        data = put(arr.tobytes() or b"\x00")
        layout = struct.pack("<BBQQ", 3, 1, data, arr.nbytes)
        return put(_object_header([(0x01, _ds_message(arr.shape)), (0x03, _dt_message(arr.dtype)), (0x08, layout)]))

    def write_group(node, node_attrs):
        entries = {}
        for name, child in node.items():
            if name == "__attrs__":
                continue
            entries[name] = write_group(child, child.get("__attrs__"))[0] if isinstance(child, dict) else write_dataset(child)
        names = sorted(entries)
        if len(names) > 2 * K_LEAF:
            raise NotImplementedError("more than 1024 entries in one group")
        heap_data, offsets = bytearray(b"\x00" * 8), {}
        for n in names:
            offsets[n] = len(heap_data)
            heap_data.extend(_pad8(n.encode() + b"\x00"))
        seg = put(bytes(heap_data) + b"\x00" * 16)
        heap = put(b"HEAP" + struct.pack("<BBBBQQQ", 0, 0, 0, 0, len(heap_data) + 16, len(heap_data), seg))
        snod = b"SNOD" + struct.pack("<BBH", 1, 0, len(names))
        for n in names:
            snod += struct.pack("<QQII", offsets[n], entries[n], 0, 0) + b"\x00" * 16
        snod_addr = put(snod + b"\x00" * (8 + 2 * K_LEAF * 40 - len(snod)))
        tree_node = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, _UNDEF, _UNDEF) + struct.pack("<QQQ", 0, snod_addr, offsets[names[-1]] if names else 0)
        btree = put(tree_node + b"\x00" * (24 + (2 * 16 + 1) * 8 + 2 * 16 * 8 - len(tree_node)))
        msgs = [(0x11, struct.pack("<QQ", btree, heap))] + [_attr_message(k, v) for k, v in (node_attrs or {}).items()]
        return put(_object_header(msgs)), btree, heap

    root, btree, heap = write_group(tree, attrs)
    sb = _SIG + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, K_LEAF, 16, 0)
    sb += struct.pack("<QQQQ", 0, _UNDEF, len(out), _UNDEF)
    sb += struct.pack("<QQII", 0, root, 1, 0) + struct.pack("<QQ", btree, heap)
    out[:len(sb)] = sb
    with open(path, "wb") as f:
        f.write(bytes(out))
    return path
