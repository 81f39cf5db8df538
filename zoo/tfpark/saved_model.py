"""TensorFlow SavedModel directories without TensorFlow: `saved_model.pb` + `variables/variables.{index,data-*}`.

Reference: `TFNet.fromSavedModel(path, tag, signature, inputs, outputs)` (zoo/.../pipeline/api/net/TFNet.scala:612-716; Python
pyzoo/zoo/pipeline/api/net/tfnet.py `from_saved_model`) loads the meta graph tagged `serve` through libtensorflow, restores the
variables and runs the graph between the signature's (or the given) tensors.  Here the same directory is READ:

* `saved_model.pb` is a `SavedModel` proto: meta graphs (tags, GraphDef, signature_def map) -- decoded with `zoo.util.protowire`;
* the variables sit in a TensorBundle: `variables.index` is a LevelDB-format table (prefix-compressed key blocks, block handles in
  an index block, a 48-byte footer) mapping a variable name to a `BundleEntryProto` (dtype, shape, shard, offset, size), and the
  bytes are in `variables.data-<shard>-of-<n>`;
* the variables are then frozen into the graph (`VariableV2` / `VarHandleOp` + their read ops become constants) and the result is
  the frozen GraphDef that `zoo.tfpark.graphdef.GraphDefModule` runs.

Fixtures of the reference that load through this file: zoo/src/test/resources/saved-model-{resource,signature}.
"""
from __future__ import annotations

import os
import struct

import numpy as np

from ..util import protowire as pw
from . import graphdef as gd

# ---- saved_model.proto / meta_graph.proto (only what is read) -----------------------------------------------------------------------
TENSOR_INFO = {1: ("name", "string"), 2: ("dtype", "enum"), 3: ("tensor_shape", "msg", gd.SHAPE)}
TENSOR_INFO_ENTRY = {1: ("key", "string"), 2: ("value", "msg", TENSOR_INFO)}
SIGNATURE = {1: ("inputs", "msg", TENSOR_INFO_ENTRY), 2: ("outputs", "msg", TENSOR_INFO_ENTRY), 3: ("method_name", "string")}
SIGNATURE_ENTRY = {1: ("key", "string"), 2: ("value", "msg", SIGNATURE)}
META_INFO = {1: ("meta_graph_version", "string"), 4: ("tags", "string"), 5: ("tensorflow_version", "string")}
META_GRAPH = {1: ("meta_info_def", "msg", META_INFO), 2: ("graph_def", "bytes"), 5: ("signature_def", "msg", SIGNATURE_ENTRY)}
SAVED_MODEL = {1: ("saved_model_schema_version", "varint"), 2: ("meta_graphs", "msg", META_GRAPH)}

# ---- tensor_bundle.proto ------------------------------------------------------------------------------------------------------------
BUNDLE_HEADER = {1: ("num_shards", "varint"), 2: ("endianness", "enum")}
BUNDLE_ENTRY = {1: ("dtype", "enum"), 2: ("shape", "msg", gd.SHAPE), 3: ("shard_id", "varint"), 4: ("offset", "varint"), 5: ("size", "varint"),
                6: ("crc32c", "fixed32")}

_TABLE_MAGIC = 0xDB4775248B80FB57


def _varint(buf, pos):
    return pw.read_varint(buf, pos)


def _snappy_uncompress(data):
    """Raw snappy block format (LevelDB tables may compress blocks; TensorFlow writes bundle indexes uncompressed)."""
    n, pos = _varint(data, 0)
    out = bytearray()
    while pos < len(data):
        tag = data[pos]
        pos += 1
        kind = tag & 3
        if kind == 0:                                   # literal
            ln = tag >> 2
            if ln >= 60:
                nb = ln - 59
                ln = int.from_bytes(data[pos:pos + nb], "little")
                pos += nb
            ln += 1
            out += data[pos:pos + ln]
            pos += ln
            continue
        if kind == 1:
            ln, off = ((tag >> 2) & 7) + 4, ((tag >> 5) << 8) | data[pos]
            pos += 1
        elif kind == 2:
            ln, off = (tag >> 2) + 1, int.from_bytes(data[pos:pos + 2], "little")
            pos += 2
        else:
            ln, off = (tag >> 2) + 1, int.from_bytes(data[pos:pos + 4], "little")
            pos += 4
        for _ in range(ln):                             # copies may overlap their own output
            out.append(out[-off])
    if len(out) != n:
        raise ValueError("snappy: length mismatch")
    return bytes(out)


def _read_block(buf, offset, size):
    """Block contents at a BlockHandle: `size` bytes + 1 byte compression type + 4 bytes crc."""
    body, ctype = buf[offset:offset + size], buf[offset + size]
    if ctype == 1:
        body = _snappy_uncompress(body)
    elif ctype != 0:
        raise NotImplementedError(f"table block compression type {ctype}")
    return body


def _block_entries(block):
    """(key, value) pairs of one table block (keys are prefix-compressed against the previous key)."""
    n_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    end = len(block) - 4 - 4 * n_restarts
    pos, key = 0, b""
    while pos < end:
        shared, pos = _varint(block, pos)
        non_shared, pos = _varint(block, pos)
        vlen, pos = _varint(block, pos)
        key = key[:shared] + bytes(block[pos:pos + non_shared])
        pos += non_shared
        yield key, bytes(block[pos:pos + vlen])
        pos += vlen


def read_table(path):
    """All (key, value) pairs of a LevelDB-format table file, in key order."""
    with open(path, "rb") as f:
        buf = f.read()
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != _TABLE_MAGIC:
        raise ValueError(f"{path} is not a table file (bad magic)")
    footer = buf[len(buf) - 48:]
    _, p = _varint(footer, 0)            # metaindex handle: offset
    _, p = _varint(footer, p)            #                   size
    ioff, p = _varint(footer, p)
    isize, p = _varint(footer, p)
    out = []
    for _, handle in _block_entries(_read_block(buf, ioff, isize)):
        boff, q = _varint(handle, 0)
        bsize, q = _varint(handle, q)
        out.extend(_block_entries(_read_block(buf, boff, bsize)))
    return out


def read_bundle(prefix):
    """TensorBundle at `<prefix>.index` / `<prefix>.data-xxxxx-of-yyyyy` -> {variable name: ndarray}."""
    entries = read_table(prefix + ".index")
    header = {"num_shards": [1]}
    tensors, shards = {}, {}
    for key, value in entries:
        if key == b"":
            header = pw.decode(value, BUNDLE_HEADER)
            if pw.one(header, "endianness", 0) != 0:
                raise NotImplementedError("big-endian tensor bundle")
            continue
        e = pw.decode(value, BUNDLE_ENTRY)
        dtype = gd.DT.get(pw.one(e, "dtype", 1))
        if dtype is None:
            continue                                     # string / variant entries (e.g. the object graph) are not weights
        shape = [int(pw.one(d, "size", 0)) for d in pw.one(e, "shape", {}).get("dim", [])]
        shard, n = int(pw.one(e, "shard_id", 0)), int(pw.one(header, "num_shards", 1))
        if shard not in shards:
            with open("%s.data-%05d-of-%05d" % (prefix, shard, n), "rb") as f:
                shards[shard] = f.read()
        off, size = int(pw.one(e, "offset", 0)), int(pw.one(e, "size", 0))
        tensors[key.decode()] = np.frombuffer(shards[shard][off:off + size], dtype=dtype).reshape(shape).copy()
    return tensors


# ---- SavedModel ---------------------------------------------------------------------------------------------------------------------
class MetaGraph:
    def __init__(self, tags, graph_def_bytes, signatures):
        self.tags, self.graph_def, self.signatures = tags, graph_def_bytes, signatures


def read_saved_model(model_path):
    """-> list of MetaGraph (tags, GraphDef bytes, {signature name: ({input key: tensor name}, {output key: tensor name})})."""
    with open(os.path.join(model_path, "saved_model.pb"), "rb") as f:
        sm = pw.decode(f.read(), SAVED_MODEL)
    out = []
    for mg in sm.get("meta_graphs", []):
        tags = [str(t) for t in pw.one(mg, "meta_info_def", {}).get("tags", [])]
        sigs = {}
        for e in mg.get("signature_def", []):
            s = pw.one(e, "value", {})
            names = lambda side: {pw.one(t, "key"): pw.one(pw.one(t, "value", {}), "name") for t in s.get(side, [])}
            sigs[pw.one(e, "key")] = (names("inputs"), names("outputs"))
        out.append(MetaGraph(tags, pw.one(mg, "graph_def", b""), sigs))
    return out


_VAR_OPS = ("VariableV2", "Variable", "VarHandleOp")


def freeze(graph_def_bytes, variables):
    """GraphDef bytes with variables -> node list where every variable is a `Const` and its read ops are identities.

    `VariableV2` (ref variables): consumers read the variable node directly or through `Identity`.  `VarHandleOp` (resource
    variables): consumers go through `ReadVariableOp`.  The bundle key of a variable is its `shared_name` attr or its node name."""
    nodes = gd.read_graphdef(graph_def_bytes)
    out = []
    for n in nodes:
        if n.op in _VAR_OPS:
            key = n.a("shared_name") or n.name
            if key not in variables and n.name in variables:
                key = n.name
            if key not in variables:
                out.append(n)                            # not restorable (e.g. an optimizer slot that was not saved); fails only if reached
                continue
            value = variables[key]
            attr = {"dtype": {"type": [gd.DT_OF[value.dtype]]},
                    "value": {"tensor": [{"dtype": [gd.DT_OF[value.dtype]], "tensor_shape": [{"dim": [{"size": [int(s)]} for s in value.shape]}],
                                          "tensor_content": [value.tobytes()]}]}}
            out.append(gd.Node(n.name, "Const", [], attr))
        elif n.op == "ReadVariableOp":
            out.append(gd.Node(n.name, "Identity", list(n.inputs), {"T": n.attr.get("dtype", {"type": [1]})}))
        else:
            out.append(n)
    return out


def load_saved_model(model_path, tag=None, signature=None, inputs=None, outputs=None):
    """-> (frozen node list, input tensor names, output tensor names).

    `tag` defaults to "serve"; input / output names default to the signature's (`signature` defaults to "serving_default", or the
    only signature); the signature's tensors are ordered by their keys."""
    tag = tag or "serve"
    metas = read_saved_model(model_path)
    meta = next((m for m in metas if tag in m.tags), None)
    if meta is None:
        raise ValueError(f"no meta graph tagged {tag!r} in {model_path} (tags: {[m.tags for m in metas]})")
    if inputs is None or outputs is None:
        if not meta.signatures:
            raise ValueError("the SavedModel has no signature: pass inputs= and outputs= tensor names")
        name = signature or ("serving_default" if "serving_default" in meta.signatures else sorted(meta.signatures)[0])
        if name not in meta.signatures:
            raise ValueError(f"signature {name!r} not found (has {sorted(meta.signatures)})")
        sin, sout = meta.signatures[name]
        inputs = inputs if inputs is not None else [sin[k] for k in sorted(sin)]
        outputs = outputs if outputs is not None else [sout[k] for k in sorted(sout)]
    prefix = os.path.join(model_path, "variables", "variables")
    variables = read_bundle(prefix) if os.path.exists(prefix + ".index") else {}
    return freeze(meta.graph_def, variables), list(inputs), list(outputs)
