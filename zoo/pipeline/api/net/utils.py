"""Helpers of the net package (reference: pyzoo/zoo/pipeline/api/net/utils.py:27-160).  `find_tensors` / `find_placeholders`
walk a graph backwards from some tensors; here the graph is a GraphDef file / bytes (read without TensorFlow).
`to_bigdl_optim_method` turns an optimizer description into the training engine's optimizer configuration."""
from __future__ import annotations

import torch


def find_tensors(sources, predicate, graph_def=None):
    """Nodes reachable backwards from `sources` (tensor / node names) for which `predicate(node)` holds, nearest first.
    `graph_def`: path, bytes or parsed node list of the graph the names refer to."""
    from ....util.tf_graph_util import _node_name, _nodes

    if graph_def is None:
        raise NotImplementedError("find_tensors(sources, predicate) walks live tf.Tensor objects, which need TensorFlow; pass the GraphDef "
                                  "the names refer to: find_tensors(names, predicate, graph_def=path_or_bytes)")
    by = {n.name: n for n in _nodes(graph_def)}
    queue, seen, out = [_node_name(s) for s in sources], set(), []
    while queue:
        name = queue.pop(0)
        if name in seen or name not in by:
            continue
        seen.add(name)
        n = by[name]
        if predicate(n):
            out.append(n)
        queue += [_node_name(i) for i in n.inputs]
    return out


def find_placeholders(grads, graph_def=None):
    return find_tensors(grads, lambda n: n.op.startswith("Placeholder"), graph_def)


def to_bigdl_optim_method(koptim_method):
    """dict of optimizers -> dict; zoo OptimMethod / torch optimizer / optimizer name -> the engine's `OptConfig`
    (the reference maps tf.keras optimizers to BigDL OptimMethods with the same hyper-parameters)."""
    from ..keras.optimizers import to_opt_config

    if isinstance(koptim_method, dict):
        return {k: to_bigdl_optim_method(v) for k, v in koptim_method.items()}
    inner = getattr(koptim_method, "optimizer", None)
    if inner is not None and not isinstance(koptim_method, torch.optim.Optimizer):
        koptim_method = inner
    return to_opt_config(koptim_method)
