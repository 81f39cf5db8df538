"""TFRecord files of `tf.train.Example` records, read and written without TensorFlow.

The reference reads them through TensorFlow (`TFDataset.from_tfrecord_file`, pyzoo/zoo/tfpark/tf_dataset.py:450-520, and
`TFRecordDataset` inside `from_tf_data_dataset`).  The container is simple enough to own: a record is
`u64 length | u32 masked_crc32c(length) | payload | u32 masked_crc32c(payload)` (the framing the TensorBoard writer of this
package already produces natively), and an Example is a protobuf message with three field kinds:

    Example   { Features features = 1; }
    Features  { map<string, Feature> feature = 1; }          // map entry: key = 1 (string), value = 2 (Feature)
    Feature   { oneof { BytesList bytes_list = 1; FloatList float_list = 2; Int64List int64_list = 3; } }
    BytesList { repeated bytes value = 1; }   FloatList { repeated float value = 1 [packed]; }   Int64List { repeated int64 value = 1 [packed]; }

Only that subset of the wire format is implemented (varint, 32-bit, length-delimited; packed and unpacked repeated scalars).
"""
from __future__ import annotations

import struct

import numpy as np

from ..native.rt import masked_crc32c, tfrecord_frame


# ------------------------------------------------------------------------------------------------ protobuf wire helpers
def _varint(n):
    n &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf, pos):
    shift = result = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7
        if shift > 70:
            raise ValueError("malformed varint")


def _ld(field, payload):
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def _fields(buf):
    """Yield (field number, wire type, value) of one message; value is int (varint / fixed) or a memoryview slice."""
    pos, n = 0, len(buf)
    while pos < n:
        key, pos = _read_varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
        elif wt == 1:
            v, pos = bytes(buf[pos:pos + 8]), pos + 8
        elif wt == 2:
            ln, pos = _read_varint(buf, pos)
            v, pos = buf[pos:pos + ln], pos + ln
        elif wt == 5:
            v, pos = bytes(buf[pos:pos + 4]), pos + 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wt}")
        yield field, wt, v


# ------------------------------------------------------------------------------------------------ tf.train.Example
def encode_example(features: dict) -> bytes:
    """{name: bytes | str | list of bytes | ndarray / list of floats | ndarray / list of ints} -> serialized tf.train.Example."""
    entries = b""
    for name in sorted(features):
        v = features[name]
        if isinstance(v, (bytes, bytearray, str)):
            v = [v]
        if isinstance(v, (list, tuple)) and v and isinstance(v[0], (bytes, bytearray, str)):
            body = b"".join(_ld(1, x.encode() if isinstance(x, str) else bytes(x)) for x in v)
            feat = _ld(1, body)
        else:
            a = np.asarray(v)
            if a.dtype.kind == "f":
                feat = _ld(2, _ld(1, a.astype("<f4").reshape(-1).tobytes()))
            elif a.dtype.kind in "iub":
                feat = _ld(3, _ld(1, b"".join(_varint(int(x)) for x in a.reshape(-1))))
            else:
                raise TypeError(f"feature {name!r}: unsupported value type {a.dtype}")
        entries += _ld(1, _ld(1, name.encode()) + _ld(2, feat))
    return _ld(1, entries)


def decode_example(data) -> dict:
    """Serialized tf.train.Example -> {name: list of bytes | float32 ndarray | int64 ndarray}."""
    out = {}
    buf = memoryview(data)
    for f, wt, features in _fields(buf):
        if f != 1 or wt != 2:
            continue
        for f2, wt2, entry in _fields(features):
            if f2 != 1 or wt2 != 2:
                continue
            name, feat = None, None
            for f3, _, v in _fields(entry):
                if f3 == 1:
                    name = bytes(v).decode()
                elif f3 == 2:
                    feat = v
            if name is None:
                continue
            out[name] = _decode_feature(feat) if feat is not None else []
    return out


def _decode_feature(feat):
    for kind, wt, body in _fields(feat):
        if wt != 2:
            continue
        if kind == 1:
            return [bytes(v) for f, _, v in _fields(body) if f == 1]
        if kind == 2:
            vals = []
            for f, w, v in _fields(body):
                if f != 1:
                    continue
                vals.append(np.frombuffer(bytes(v), "<f4") if w == 2 else np.frombuffer(v, "<f4"))      # packed | single fixed32
            return np.concatenate(vals).astype(np.float32) if vals else np.zeros(0, np.float32)
        if kind == 3:
            vals = []
            for f, w, v in _fields(body):
                if f != 1:
                    continue
                if w == 2:
                    pos, n = 0, len(v)
                    while pos < n:
                        x, pos = _read_varint(v, pos)
                        vals.append(x)
                else:
                    vals.append(v)
            a = np.asarray(vals, dtype=np.uint64).astype(np.int64) if vals else np.zeros(0, np.int64)     # two's complement
            return a
    return []


# ------------------------------------------------------------------------------------------------ record files
def write_tfrecord(path, records, append=False):
    """Write serialized records (bytes) or feature dicts (encoded as tf.train.Example) to a TFRecord file; returns the count."""
    n = 0
    with open(path, "ab" if append else "wb") as fh:
        for r in records:
            fh.write(tfrecord_frame(encode_example(r) if isinstance(r, dict) else bytes(r)))
            n += 1
    return n


def read_tfrecord(path, check_crc=True):
    """Iterate the payloads of a TFRecord file; a truncated tail or a checksum mismatch raises IOError."""
    with open(path, "rb") as fh:
        while True:
            head = fh.read(12)
            if not head:
                return
            if len(head) < 12:
                raise IOError(f"{path}: truncated record header")
            (length,), (crc_len,) = struct.unpack("<Q", head[:8]), struct.unpack("<I", head[8:])
            if check_crc and masked_crc32c(head[:8]) != crc_len:
                raise IOError(f"{path}: corrupted record length")
            body = fh.read(length + 4)
            if len(body) < length + 4:
                raise IOError(f"{path}: truncated record")
            payload = body[:length]
            if check_crc and masked_crc32c(payload) != struct.unpack("<I", body[length:])[0]:
                raise IOError(f"{path}: corrupted record payload")
            yield payload


def read_examples(paths, check_crc=True):
    """Decoded tf.train.Example dicts of one file, a glob pattern, or a list of files."""
    import glob as _glob

    if isinstance(paths, str):
        paths = sorted(_glob.glob(paths)) or [paths]
    for p in paths:
        for payload in read_tfrecord(p, check_crc):
            yield decode_example(payload)


_IMAGE_MAGIC = (b"\x89PNG", b"\xff\xd8\xff", b"BM")


def _auto_decode(b):
    if isinstance(b, (bytes, bytearray)) and bytes(b[:4]).startswith(_IMAGE_MAGIC):
        import cv2

        img = cv2.imdecode(np.frombuffer(b, np.uint8), cv2.IMREAD_UNCHANGED)
        if img is not None:
            return img
    return b


def is_numeric_feature(value):
    """False for bytes features that are not encoded images (they cannot become tensors)."""
    if isinstance(value, list):
        return all(not isinstance(_auto_decode(b), (bytes, bytearray)) for b in value)
    return True


def examples_to_arrays(examples, features, labels=None, image_decoder=None):
    """Stack Example features into ndarrays.  `features` / `labels`: a name or a list of names.  Bytes features: `image_decoder`
    (a callable bytes -> array) decodes them; with `image_decoder=None` bytes that start with a PNG / JPEG / BMP signature are decoded
    with cv2 (-> HW or HWC uint8) and other bytes (file names, format tags) stay an object array."""
    names_x = [features] if isinstance(features, str) else list(features)
    names_y = [] if labels is None else ([labels] if isinstance(labels, str) else list(labels))
    cols = {n: [] for n in names_x + names_y}
    for ex in examples:
        for n in cols:
            v = ex[n]
            if isinstance(v, list):                        # bytes feature
                v = [image_decoder(b) for b in v] if image_decoder is not None else [_auto_decode(b) for b in v]
                v = v[0] if len(v) == 1 else v
            cols[n].append(v)

    def stack(n):
        vals = cols[n]
        try:
            return np.stack([np.asarray(v) for v in vals])
        except ValueError:
            return np.asarray(vals, dtype=object)

    x = [stack(n) for n in names_x]
    y = [stack(n) for n in names_y]
    return (x[0] if len(x) == 1 else x), (None if not y else (y[0] if len(y) == 1 else y))
