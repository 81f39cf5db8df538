"""Schema-light protobuf reading and writing (binary wire format and text format), without generated classes.

The foreign-model importers (Caffe `NetParameter`, TensorFlow `GraphDef`) and their tests need to read -- and, for tests and
exporters, write -- protobuf files whose `.proto` definitions are not shipped with this framework.  A message is represented
as `Msg`: an ordered multimap field-number -> raw values; a *schema* (a plain dict) gives names and types to the field
numbers one cares about, everything else is skipped.  Text format (Caffe `.prototxt`, `pbtxt`) parses into nested dicts of
lists.

Schema:  {field_number: (name, kind[, sub_schema])} with kind in
    "varint" | "sint" | "bool" | "enum" | "fixed32" | "fixed64" | "float" | "double" | "string" | "bytes" | "msg" |
    "packed_float" | "packed_varint" | "packed_double"        (packed kinds also accept the unpacked encoding)
`decode(buf, schema)` returns {name: [values...]} (every field is a list; use `one()` for singular fields).
"""
from __future__ import annotations

import re
import struct

VARINT, FIXED64, LEN, FIXED32 = 0, 1, 2, 5


def write_varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def read_varint(buf, pos):
    shift, val = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7
        if shift > 70:
            raise ValueError("malformed varint")


def iter_fields(buf):
    """Yield (field_number, wire_type, value) -- value is an int (varint / fixed) or a memoryview (length-delimited)."""
    mv = memoryview(buf)
    pos, n = 0, len(mv)
    while pos < n:
        key, pos = read_varint(mv, pos)
        field, wt = key >> 3, key & 7
        if wt == VARINT:
            v, pos = read_varint(mv, pos)
        elif wt == FIXED64:
            v = struct.unpack_from("<Q", mv, pos)[0]
            pos += 8
        elif wt == LEN:
            ln, pos = read_varint(mv, pos)
            if pos + ln > n:
                raise ValueError("truncated length-delimited field")
            v = mv[pos:pos + ln]
            pos += ln
        elif wt == FIXED32:
            v = struct.unpack_from("<I", mv, pos)[0]
            pos += 4
        else:
            raise ValueError(f"unsupported wire type {wt} (groups are not supported)")
        yield field, wt, v


def _signed(v, bits=64):
    return v - (1 << bits) if v >> (bits - 1) else v


def decode(buf, schema):
    out = {}
    for field, wt, v in iter_fields(buf):
        ent = schema.get(field)
        if ent is None:
            continue
        name, kind = ent[0], ent[1]
        dst = out.setdefault(name, [])
        if kind == "msg":
            dst.append(decode(v, ent[2]))
        elif kind == "string":
            dst.append(bytes(v).decode("utf-8", errors="replace"))
        elif kind == "bytes":
            dst.append(bytes(v))
        elif kind in ("varint", "enum"):
            dst.append(_signed(v) if kind == "varint" else v)
        elif kind == "sint":
            dst.append((v >> 1) ^ -(v & 1))
        elif kind == "bool":
            dst.append(bool(v))
        elif kind == "float":
            dst.append(struct.unpack("<f", struct.pack("<I", v))[0])
        elif kind == "double":
            dst.append(struct.unpack("<d", struct.pack("<Q", v))[0])
        elif kind in ("fixed32", "fixed64"):
            dst.append(v)
        elif kind == "packed_float":
            if wt == LEN:
                import numpy as np

                dst.append(np.frombuffer(v, dtype="<f4"))
            else:
                dst.append(struct.unpack("<f", struct.pack("<I", v))[0])
        elif kind == "packed_double":
            if wt == LEN:
                import numpy as np

                dst.append(np.frombuffer(v, dtype="<f8"))
            else:
                dst.append(struct.unpack("<d", struct.pack("<Q", v))[0])
        elif kind == "packed_varint":
            if wt == LEN:
                pos, vals = 0, []
                while pos < len(v):
                    x, pos = read_varint(v, pos)
                    vals.append(_signed(x))
                dst.extend(vals)
            else:
                dst.append(_signed(v))
        else:
            raise ValueError(f"unknown schema kind {kind!r}")
    return out


def one(msg, name, default=None):
    v = msg.get(name)
    return v[-1] if v else default


def floats(msg, name):
    """Concatenate a repeated/packed float field into one numpy array."""
    import numpy as np

    parts = msg.get(name) or []
    if not parts:
        return np.zeros(0, dtype=np.float32)
    return np.concatenate([np.atleast_1d(np.asarray(p, dtype=np.float32)) for p in parts])


# ------------------------------------------------------------------------------------------------------------ writing --
def _key(field, wt):
    return write_varint((field << 3) | wt)


def encode(values, schema):
    """Inverse of `decode`: values = {name: value or [values]}; field order follows the schema."""
    by_name = {ent[0]: (num, ent) for num, ent in schema.items()}
    out = bytearray()
    for name, val in values.items():
        if name not in by_name:
            raise KeyError(f"field {name!r} is not in the schema")
        num, ent = by_name[name]
        kind = ent[1]
        vals = val if isinstance(val, (list, tuple)) and kind not in ("packed_float", "packed_varint", "packed_double") else [val]
        for v in vals:
            if kind == "msg":
                payload = encode(v, ent[2])
                out += _key(num, LEN) + write_varint(len(payload)) + payload
            elif kind == "string":
                b = v.encode("utf-8")
                out += _key(num, LEN) + write_varint(len(b)) + b
            elif kind == "bytes":
                out += _key(num, LEN) + write_varint(len(v)) + bytes(v)
            elif kind in ("varint", "enum", "bool"):
                out += _key(num, VARINT) + write_varint(int(v))
            elif kind == "sint":
                out += _key(num, VARINT) + write_varint((int(v) << 1) ^ (int(v) >> 63))
            elif kind == "float":
                out += _key(num, FIXED32) + struct.pack("<f", float(v))
            elif kind == "double":
                out += _key(num, FIXED64) + struct.pack("<d", float(v))
            elif kind == "fixed32":
                out += _key(num, FIXED32) + struct.pack("<I", int(v))
            elif kind == "fixed64":
                out += _key(num, FIXED64) + struct.pack("<Q", int(v))
            elif kind == "packed_float":
                import numpy as np

                b = np.asarray(v, dtype="<f4").tobytes()
                out += _key(num, LEN) + write_varint(len(b)) + b
            elif kind == "packed_double":
                import numpy as np

                b = np.asarray(v, dtype="<f8").tobytes()
                out += _key(num, LEN) + write_varint(len(b)) + b
            elif kind == "packed_varint":
                b = b"".join(write_varint(int(x)) for x in v)
                out += _key(num, LEN) + write_varint(len(b)) + b
            else:
                raise ValueError(f"unknown schema kind {kind!r}")
    return bytes(out)


# -------------------------------------------------------------------------------------------------------- text format --
_TOKEN = re.compile(r"""\s*(?:(\#[^\n]*)|([{}<>:;,\[\]])|("(?:\\.|[^"\\])*"|'(?:\\.|[^'\\])*')|([^\s{}<>:;,\[\]"'#]+))""")


def _tokens(text):
    pos, n = 0, len(text)
    while pos < n:
        m = _TOKEN.match(text, pos)
        if not m:
            if text[pos:].strip() == "":
                return
            raise ValueError(f"text format: cannot tokenize at offset {pos}: {text[pos:pos + 30]!r}")
        pos = m.end()
        if m.group(1):
            continue
        if m.group(2):
            yield ("p", m.group(2))
        elif m.group(3):
            yield ("s", _unescape(m.group(3)[1:-1]))
        else:
            yield ("w", m.group(4))


def _unescape(s):
    return bytes(s, "utf-8").decode("unicode_escape") if "\\" in s else s


def _scalar(kind, tok):
    if kind == "s":
        return tok
    if tok in ("true", "True"):
        return True
    if tok in ("false", "False"):
        return False
    try:
        return int(tok, 0)
    except ValueError:
        pass
    try:
        return float(tok.rstrip("f"))
    except ValueError:
        return tok       # an enum identifier


def parse_text(text):
    """Protobuf text format -> {field: [values]}; nested messages are dicts of the same shape."""
    toks = list(_tokens(text))
    pos = 0

    def message(closer):
        nonlocal pos
        out = {}
        while pos < len(toks):
            kind, tok = toks[pos]
            if kind == "p" and tok in ("}", ">"):
                if closer is None:
                    raise ValueError("text format: unbalanced closing brace")
                pos += 1
                return out
            if kind == "p" and tok in (";", ","):
                pos += 1
                continue
            if kind != "w":
                raise ValueError(f"text format: expected a field name, got {tok!r}")
            name = tok
            pos += 1
            if pos < len(toks) and toks[pos] == ("p", ":"):
                pos += 1
            k2, t2 = toks[pos]
            if k2 == "p" and t2 in ("{", "<"):
                pos += 1
                out.setdefault(name, []).append(message(t2))
            elif k2 == "p" and t2 == "[":
                pos += 1
                while toks[pos] != ("p", "]"):
                    if toks[pos] == ("p", ","):
                        pos += 1
                        continue
                    if toks[pos][0] == "p" and toks[pos][1] in ("{", "<"):
                        pos += 1
                        out.setdefault(name, []).append(message("{"))
                    else:
                        out.setdefault(name, []).append(_scalar(*toks[pos]))
                        pos += 1
                pos += 1
            else:
                val = _scalar(k2, t2)
                pos += 1
                while k2 == "s" and pos < len(toks) and toks[pos][0] == "s":      # adjacent string literals concatenate
                    val += toks[pos][1]
                    pos += 1
                out.setdefault(name, []).append(val)
        if closer is not None:
            raise ValueError("text format: missing closing brace")
        return out

    return message(None)
