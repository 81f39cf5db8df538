"""Hadoop SequenceFiles of images without Hadoop.

The reference keeps ImageNet as SequenceFiles (`Text` key = "name\\nlabel" or "label", `Text` value = 4-byte width, 4-byte height,
then the raw BGR pixels) written by BigDL's ImageNetSeqFileGenerator and read through `sc.sequenceFile`:
`ImageSet.readSequenceFiles(path, sc, minPartitions, classNum)` (zoo/.../feature/image/ImageSet.scala:333-352) and the
Inception / ResNet examples (examples/inception/ImageNet2012.scala:45-70, examples/resnet/Utils.scala:118-140).

`read_sequence_file` decodes the container format itself (version 6: header with key / value class names, compression flags,
metadata, 16-byte sync marker; records as length-prefixed key / value, `-1` escapes before a sync marker), uncompressed,
record-compressed or block-compressed with the default (zlib) codec.  `write_sequence_file` writes the uncompressed form.
Fixture read in the tests: zoo/src/test/resources/pmem/mini_imagenet_seq.
"""
from __future__ import annotations

import glob
import os
import struct
import zlib

import numpy as np

_TEXT = "org.apache.hadoop.io.Text"
_BYTES = "org.apache.hadoop.io.BytesWritable"
_DEFAULT_CODEC = "org.apache.hadoop.io.compress.DefaultCodec"


def _read_vint(buf, pos):
    """Hadoop WritableUtils.readVInt/readVLong."""
    first = struct.unpack_from("b", buf, pos)[0]
    pos += 1
    if first >= -112:
        return first, pos
    size = (-119 - first) if first < -120 else (-111 - first)       # total bytes including the first
    val = int.from_bytes(buf[pos:pos + size - 1], "big")
    pos += size - 1
    return (~val if first < -120 else val), pos


def _write_vint(v):
    if -112 <= v <= 127:
        return struct.pack("b", v)
    neg = v < 0
    if neg:
        v = ~v
    n = (v.bit_length() + 7) // 8
    return struct.pack("b", (-120 - n) if neg else (-112 - n)) + v.to_bytes(n, "big")


def _read_string(buf, pos):
    n, pos = _read_vint(buf, pos)
    return bytes(buf[pos:pos + n]).decode(), pos + n


def _payload(cls, raw):
    """The bytes a Writable of class `cls` carries (Text: vint length + bytes; BytesWritable: int length + bytes)."""
    if cls == _TEXT:
        n, p = _read_vint(raw, 0)
        return bytes(raw[p:p + n])
    if cls == _BYTES:
        n = struct.unpack_from(">i", raw, 0)[0]
        return bytes(raw[4:4 + n])
    return bytes(raw)


def read_sequence_file(path):
    """Yields (key bytes, value bytes) of one SequenceFile."""
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:3] != b"SEQ":
        raise ValueError(f"{path} is not a SequenceFile")
    version = buf[3]
    if version < 5:
        raise NotImplementedError(f"SequenceFile version {version}")
    pos = 4
    kcls, pos = _read_string(buf, pos)
    vcls, pos = _read_string(buf, pos)
    compressed, block = bool(buf[pos]), bool(buf[pos + 1])
    pos += 2
    if compressed:
        codec, pos = _read_string(buf, pos)
        if codec != _DEFAULT_CODEC:
            raise NotImplementedError(f"SequenceFile codec {codec} (only the default zlib codec is read)")
    n_meta = struct.unpack_from(">i", buf, pos)[0]
    pos += 4
    for _ in range(n_meta):
        _, pos = _read_string(buf, pos)
        _, pos = _read_string(buf, pos)
    sync = bytes(buf[pos:pos + 16])
    pos += 16
    end = len(buf)
    if block:
        while pos < end:
            if struct.unpack_from(">i", buf, pos)[0] != -1 or bytes(buf[pos + 4:pos + 20]) != sync:
                raise ValueError("block-compressed SequenceFile: sync marker expected")
            pos += 20
            n, pos = _read_vint(buf, pos)
            parts = []
            for _ in range(4):                                     # key lengths, keys, value lengths, values
                size, pos = _read_vint(buf, pos)
                parts.append(zlib.decompress(bytes(buf[pos:pos + size])))
                pos += size
            kl, ks, vl, vs = parts
            kp = vp = klp = vlp = 0
            for _ in range(n):
                klen, klp = _read_vint(kl, klp)
                vlen, vlp = _read_vint(vl, vlp)
                yield _payload(kcls, ks[kp:kp + klen]), _payload(vcls, vs[vp:vp + vlen])
                kp, vp = kp + klen, vp + vlen
        return
    while pos < end:
        rec_len = struct.unpack_from(">i", buf, pos)[0]
        pos += 4
        if rec_len == -1:                                          # sync escape
            if bytes(buf[pos:pos + 16]) != sync:
                raise ValueError("SequenceFile: bad sync marker")
            pos += 16
            continue
        key_len = struct.unpack_from(">i", buf, pos)[0]
        pos += 4
        key, val = buf[pos:pos + key_len], buf[pos + key_len:pos + rec_len]
        pos += rec_len
        if compressed:
            val = zlib.decompress(bytes(val))
        yield _payload(kcls, key), _payload(vcls, val)


def write_sequence_file(path, records, sync_every=2000):
    """records: iterable of (key bytes, value bytes) -> an uncompressed Text / Text SequenceFile (what the reference's generators write)."""
    sync = os.urandom(16)
    with open(path, "wb") as f:
        f.write(b"SEQ\x06" + _write_vint(len(_TEXT)) + _TEXT.encode() + _write_vint(len(_TEXT)) + _TEXT.encode() + b"\x00\x00" +
                struct.pack(">i", 0) + sync)
        since = 0
        for key, val in records:
            k = _write_vint(len(key)) + bytes(key)
            v = _write_vint(len(val)) + bytes(val)
            if since >= sync_every:
                f.write(struct.pack(">i", -1) + sync)
                since = 0
            f.write(struct.pack(">ii", len(k) + len(v), len(k)) + k + v)
            since += len(k) + len(v) + 8
    return path


def read_label(key_text):
    """Label from a key: "name\\nlabel" or just "label" (ImageNet2012.scala:62-69)."""
    parts = key_text.split("\n")
    return parts[0] if len(parts) == 1 else parts[1]


def decode_image_record(value):
    """Value bytes -> (H, W, 3) uint8 BGR mat: two big-endian ints (width, height) then the pixels."""
    w, h = struct.unpack_from(">ii", value, 0)
    return np.frombuffer(value, np.uint8, count=3 * w * h, offset=8).reshape(h, w, 3)


def encode_image_record(mat):
    mat = np.ascontiguousarray(mat, dtype=np.uint8)
    return struct.pack(">ii", mat.shape[1], mat.shape[0]) + mat.tobytes()


def sequence_file_paths(path):
    """A file, or the part files of a folder (hidden files and _SUCCESS are skipped, as Hadoop's input formats do)."""
    if os.path.isdir(path):
        return sorted(p for p in glob.glob(os.path.join(path, "*")) if os.path.isfile(p) and not os.path.basename(p).startswith(("_", ".")))
    return sorted(glob.glob(path)) or [path]


def read_image_records(path, class_num=1000):
    """Yields (name or None, label float, mat) for every image of the SequenceFiles under `path` with label <= class_num."""
    for p in sequence_file_paths(path):
        for key, value in read_sequence_file(p):
            text = key.decode()
            label = float(read_label(text))
            if label <= class_num:
                yield (text.split("\n")[0] if "\n" in text else None), label, decode_image_record(value)
