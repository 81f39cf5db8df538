"""TensorBoard event-file writer / reader without TensorFlow.

Reference: zoo/src/main/scala/.../tensorboard/{Summary,FileWriter,EventWriter,RecordWriter,
FileReader}.scala (553 LoC): `Summary.addScalar/addHistogram`, an `EventWriter` thread draining a
queue into TFRecord-framed `Event` protobufs with masked CRC32C, and `FileReader.readScalar`
returning `(step, value, wallclock)` triples that `get_train_summary(tag)` hands to the user
(Topology.scala:222-245).  The protobuf messages are tiny, so they are encoded by hand.
CRC32C comes from the C++ runtime when it is built, else from a table in Python.
"""
from __future__ import annotations

import os
import queue
import socket
import struct
import threading
import time

import numpy as np

_CRC_TABLE = None


def _crc_table():
    global _CRC_TABLE
    if _CRC_TABLE is None:
        t = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            t.append(c)
        _CRC_TABLE = t
    return _CRC_TABLE


def crc32c(data: bytes) -> int:
    try:
        from ..native.runtime import crc32c as native_crc

        return native_crc(data)
    except Exception:
        pass
    t, c = _crc_table(), 0xFFFFFFFF
    for b in data:
        c = t[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc(data: bytes) -> int:
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


# ----------------------------------------------------------- protobuf encoding --
def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(field, wire):
    return _varint((field << 3) | wire)


def _f_double(field, v):
    return _key(field, 1) + struct.pack("<d", v)


def _f_float(field, v):
    return _key(field, 5) + struct.pack("<f", v)


def _f_varint(field, v):
    return _key(field, 0) + _varint(v)


def _f_bytes(field, b):
    return _key(field, 2) + _varint(len(b)) + b


def _packed_doubles(field, vals):
    return _f_bytes(field, b"".join(struct.pack("<d", float(v)) for v in vals))


def encode_scalar_event(tag, value, step, wall_time=None):
    val = _f_bytes(1, tag.encode()) + _f_float(2, float(value))              # Summary.Value{tag, simple_value}
    summ = _f_bytes(1, val)                                                   # Summary{value}
    return _f_double(1, wall_time or time.time()) + _f_varint(2, int(step)) + _f_bytes(5, summ)


def encode_histogram_event(tag, values, step, wall_time=None):
    v = np.asarray(values, dtype=np.float64).ravel()
    if v.size == 0:
        v = np.zeros(1)
    counts, edges = np.histogram(v, bins=30)
    histo = (_f_double(1, v.min()) + _f_double(2, v.max()) + _f_double(3, v.size) + _f_double(4, v.sum()) +
             _f_double(5, float((v * v).sum())) + _packed_doubles(6, edges[1:]) + _packed_doubles(7, counts))
    val = _f_bytes(1, tag.encode()) + _f_bytes(5, histo)
    return _f_double(1, wall_time or time.time()) + _f_varint(2, int(step)) + _f_bytes(5, _f_bytes(1, val))


def encode_file_version_event():
    return _f_double(1, time.time()) + _f_bytes(3, b"brain.Event:2")


def frame_record(data: bytes) -> bytes:
    header = struct.pack("<Q", len(data))
    return header + struct.pack("<I", masked_crc(header)) + data + struct.pack("<I", masked_crc(data))


# -------------------------------------------------------------------- writer --
class FileWriter:
    """Asynchronous event-file writer (EventWriter thread + RecordWriter of the reference)."""

    def __init__(self, log_dir, flush_secs=2.0):
        os.makedirs(log_dir, exist_ok=True)
        self.log_dir = log_dir
        self.path = os.path.join(log_dir, f"events.out.tfevents.{int(time.time())}.{socket.gethostname()}.{os.getpid()}")
        self._f = open(self.path, "ab")
        self._q = queue.Queue()
        self._flush_secs = flush_secs
        self._closed = False
        self._f.write(frame_record(encode_file_version_event()))
        self._t = threading.Thread(target=self._run, daemon=True, name="zoo-event-writer")
        self._t.start()

    def _run(self):
        last = time.time()
        while True:
            try:
                item = self._q.get(timeout=self._flush_secs)
            except queue.Empty:
                item = b""
            if item is None:
                break
            if item:
                self._f.write(frame_record(item))
            if time.time() - last >= self._flush_secs or self._q.empty():
                self._f.flush()
                last = time.time()
        self._f.flush()

    def add_event(self, payload: bytes):
        if not self._closed:
            self._q.put(payload)
        return self

    def flush(self):
        if self._closed:          # everything was flushed by close(); reading back afterwards is fine
            return
        deadline = time.time() + 5
        while not self._q.empty() and time.time() < deadline:
            time.sleep(0.005)
        time.sleep(0.005)
        self._f.flush()

    def close(self):
        if self._closed:
            return
        self._closed = True
        self._q.put(None)
        self._t.join(timeout=5)
        self._f.close()


class FileReader:
    """`readScalar(folder, tag)` -> [(step, value, wall_time)] (FileReader.scala)."""

    @staticmethod
    def list_events(folder):
        return sorted(os.path.join(folder, f) for f in os.listdir(folder) if "tfevents" in f) if os.path.isdir(folder) else []

    @staticmethod
    def _records(path):
        with open(path, "rb") as f:
            while True:
                h = f.read(8)
                if len(h) < 8:
                    return
                (n,) = struct.unpack("<Q", h)
                crc_h = f.read(4)
                data = f.read(n)
                crc_d = f.read(4)
                if len(data) < n or len(crc_d) < 4:
                    return
                if struct.unpack("<I", crc_h)[0] != masked_crc(h) or struct.unpack("<I", crc_d)[0] != masked_crc(data):
                    raise IOError(f"corrupt TFRecord in {path}")
                yield data

    @staticmethod
    def _parse(buf):
        """Minimal protobuf walker: yields (field, wire, value)."""
        i, n = 0, len(buf)
        while i < n:
            k = 0
            s = 0
            while True:
                b = buf[i]; i += 1
                k |= (b & 0x7F) << s
                s += 7
                if not b & 0x80:
                    break
            field, wire = k >> 3, k & 7
            if wire == 0:
                v = 0; s = 0
                while True:
                    b = buf[i]; i += 1
                    v |= (b & 0x7F) << s
                    s += 7
                    if not b & 0x80:
                        break
            elif wire == 1:
                v = buf[i:i + 8]; i += 8
            elif wire == 5:
                v = buf[i:i + 4]; i += 4
            elif wire == 2:
                ln = 0; s = 0
                while True:
                    b = buf[i]; i += 1
                    ln |= (b & 0x7F) << s
                    s += 7
                    if not b & 0x80:
                        break
                v = buf[i:i + ln]; i += ln
            else:
                return
            yield field, wire, v

    @classmethod
    def read_scalar(cls, folder, tag):
        out = []
        for path in cls.list_events(folder):
            for rec in cls._records(path):
                wall, step, summ = 0.0, 0, None
                for f, w, v in cls._parse(rec):
                    if f == 1 and w == 1:
                        wall = struct.unpack("<d", v)[0]
                    elif f == 2 and w == 0:
                        step = v
                    elif f == 5 and w == 2:
                        summ = v
                if summ is None:
                    continue
                for f, w, val in cls._parse(summ):
                    if f != 1 or w != 2:
                        continue
                    t, sv = None, None
                    for ff, ww, vv in cls._parse(val):
                        if ff == 1 and ww == 2:
                            t = bytes(vv).decode()
                        elif ff == 2 and ww == 5:
                            sv = struct.unpack("<f", vv)[0]
                    if t == tag and sv is not None:
                        out.append((step, sv, wall))
        out.sort(key=lambda r: r[0])
        return out

    readScalar = read_scalar


class Summary:
    """`Summary(log_dir, app_name)` writing to `<log_dir>/<app_name>/<kind>`."""

    kind = ""

    def __init__(self, log_dir, app_name):
        self.log_dir, self.app_name = log_dir, app_name
        self.folder = os.path.join(log_dir, app_name, self.kind) if self.kind else os.path.join(log_dir, app_name)
        self.writer = FileWriter(self.folder)
        self._triggers = {}

    def add_scalar(self, tag, value, step):
        self.writer.add_event(encode_scalar_event(tag, value, step))
        return self

    def add_histogram(self, tag, values, step):
        self.writer.add_event(encode_histogram_event(tag, values, step))
        return self

    addScalar, addHistogram = add_scalar, add_histogram

    def read_scalar(self, tag):
        self.writer.flush()
        return FileReader.read_scalar(self.folder, tag)

    readScalar = read_scalar

    def set_summary_trigger(self, name, trigger):
        """Per-tag trigger (e.g. 'Parameters' every N iterations), TrainSummary.setSummaryTrigger."""
        self._triggers[name] = trigger
        return self

    def get_summary_trigger(self, name):
        return self._triggers.get(name)

    def close(self):
        self.writer.close()


class TrainSummary(Summary):
    kind = "train"


class ValidationSummary(Summary):
    kind = "validation"


class InferenceSummary(Summary):
    """Serving throughput summary (zoo/.../pipeline/inference/InferenceSummary.scala:24-40)."""
    kind = "inference"
