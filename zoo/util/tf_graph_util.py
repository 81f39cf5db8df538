"""GraphDef utilities (reference: pyzoo/zoo/util/tf_graph_util.py:1-360, which carries TensorFlow's graph_util_impl).  The
reference versions work on `tf.GraphDef` objects; these work on GraphDef FILES / BYTES through the framework's own protobuf reader
(`zoo.tfpark.graphdef`), so they need no TensorFlow.  A "graph_def" argument is a path, the serialized bytes, or a list of
parsed nodes; functions that return a graph return its serialized bytes.
"""
from __future__ import annotations

import numpy as np

from ..tfpark.graphdef import DT, Node, read_graphdef, write_graphdef
from .tf import node_tuple


def _nodes(graph_def):
    if isinstance(graph_def, (list, tuple)) and (not graph_def or isinstance(graph_def[0], Node)):
        return list(graph_def)
    return read_graphdef(graph_def)


def _node_name(n):
    return n.split(":")[0].lstrip("^")


def must_run_on_cpu(node, pin_variables_on_cpu=False):
    """True for the ops TensorFlow only has host kernels for: string / int32 constants, string casts, and (optionally) variables."""
    if pin_variables_on_cpu and node.op in ("Variable", "VariableV2", "VarHandleOp", "AutoReloadVariable", "MutableHashTable"):
        return True
    if node.op == "Const":
        return node.a("dtype") in (3, 7)          # DT_INT32, DT_STRING
    if node.op in ("DynamicStitch", "ParallelDynamicStitch"):
        return node.a("T") in (3, 7)
    if node.op == "Cast":
        return node.a("SrcT") == 7
    return False


def _bfs(nodes, dest_nodes):
    by = {n.name: n for n in nodes}
    for d in dest_nodes:
        if _node_name(d) not in by:
            raise AssertionError(f"{d} is not in graph")
    seen, stack = set(), [_node_name(d) for d in dest_nodes]
    while stack:
        name = stack.pop()
        if name in seen:
            continue
        seen.add(name)
        stack += [_node_name(i) for i in by[name].inputs if _node_name(i) in by]
    return seen


def extract_sub_graph(graph_def, dest_nodes):
    """The sub-graph that reaches `dest_nodes` (node names), in the original node order."""
    if isinstance(dest_nodes, str):
        raise TypeError("dest_nodes must be a list.")
    nodes = _nodes(graph_def)
    keep = _bfs(nodes, dest_nodes)
    return write_graphdef(None, [node_tuple(n) for n in nodes if n.name in keep])


def tensor_shape_from_node_def_name(graph, input_name):
    """Static shape of a node's first output where the GraphDef records one (constants, placeholders with a shape attribute,
    `_output_shapes`); None otherwise."""
    by = {n.name: n for n in _nodes(graph)}
    n = by[_node_name(input_name)]
    if n.op == "Const":
        return list(np.asarray(n.a("value")).shape)
    for key in ("shape", "_output_shapes"):
        v = n.a(key)
        if v is not None:
            return list(v[0]) if key == "_output_shapes" and v and isinstance(v[0], (list, tuple)) else list(v)
    return None


def convert_variables_to_constants(sess, input_graph_def, output_node_names, variable_names_whitelist=None, variable_names_blacklist=None):
    """Freeze variables into constants.  `sess`: a mapping variable name -> ndarray (the values a tf.Session would have supplied).
    Variable nodes (`Variable`, `VariableV2`, `VarHandleOp` + its `ReadVariableOp`) become `Const`; the result is pruned to
    `output_node_names`."""
    values = dict(sess or {})
    nodes = _nodes(input_graph_def)
    keep = _bfs(nodes, output_node_names)
    out, frozen = [], set()
    for n in nodes:
        if n.name not in keep:
            continue
        if n.op in ("Variable", "VariableV2", "VarHandleOp"):
            ok = (variable_names_whitelist is None or n.name in variable_names_whitelist) and \
                 (variable_names_blacklist is None or n.name not in variable_names_blacklist)
            if ok:
                if n.name not in values:
                    raise KeyError(f"no value given for variable {n.name}")
                arr = np.ascontiguousarray(values[n.name])
                out.append((n.name, "Const", [], {"dtype": arr.dtype.type, "value": arr}))
                frozen.add(n.name)
                continue
        if n.op == "ReadVariableOp" and _node_name(n.inputs[0]) in frozen:
            out.append((n.name, "Identity", [n.inputs[0]], {"T": DT.get(n.a("dtype"), np.float32)}))
            continue
        out.append(node_tuple(n))
    return write_graphdef(None, out)


def remove_training_nodes(input_graph, protected_nodes=None):
    """Drop `CheckNumerics` nodes and splice out `Identity` / `StopGradient` pass-throughs that are not protected and carry no
    control edges, rewiring their consumers to their input."""
    protected = set(protected_nodes or [])
    nodes = _nodes(input_graph)
    splice = {}
    for n in nodes:
        if n.name in protected:
            continue
        if n.op == "CheckNumerics" or (n.op in ("Identity", "StopGradient") and not any(i.startswith("^") for i in n.inputs)):
            splice[n.name] = n.inputs[0]
    controlled = {_node_name(i) for n in nodes for i in n.inputs if i.startswith("^")}
    for name in list(splice):
        if name in controlled:           # something waits on it through a control edge: keep it
            del splice[name]

    def resolve(ref):
        ctrl = ref.startswith("^")
        base = _node_name(ref)
        hops = 0
        while base in splice and hops < len(nodes):
            ref = splice[base]
            base = _node_name(ref)
            hops += 1
        return ("^" + base) if ctrl else ref

    out = []
    for n in nodes:
        if n.name in splice:
            continue
        name, op, ins, attrs = node_tuple(n)
        out.append((name, op, [resolve(i) for i in ins], attrs))
    return write_graphdef(None, out)
