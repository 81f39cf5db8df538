"""GraphNet: graph surgery on a loaded model (reference: pyzoo/zoo/pipeline/api/net/graph_net.py:32-150,
NetUtils.scala GraphNet): `new_graph(outputs)`, `freeze_up_to(names)`, `unfreeze`, `to_keras`, `layers`."""
from __future__ import annotations

import torch.nn as nn

from ..keras.engine.topology import KerasNet, Model


class GraphNet(KerasNet):
    def __init__(self, model: nn.Module):
        super().__init__()
        self.model = model

    def forward(self, *xs):
        return self.model(*xs)

    def _named(self):
        out = []
        for n, m in self.model.named_modules():
            if n:
                out.append((getattr(m, "zoo_name", None) or n, n, m))
        return out

    def new_graph(self, outputs):
        """A model that stops at the named layer(s).  Functional `Model`s cut the DAG; any other module is cut with
        a forward hook at the named sub-module."""
        if isinstance(self.model, Model) and self.model.graph is not None:
            return GraphNet(self.model.new_graph(outputs))
        names = [outputs] if isinstance(outputs, str) else list(outputs)
        table = {a: m for a, _, m in self._named()} | {b: m for _, b, m in self._named()}
        missing = [n for n in names if n not in table]
        if missing:
            raise KeyError(f"no layer named {missing}")
        return GraphNet(_Truncated(self.model, [table[n] for n in names]))

    def freeze_up_to(self, names):
        names = [names] if isinstance(names, str) else list(names)
        if isinstance(self.model, Model) and self.model.graph is not None:
            self.model.freeze_up_to(names)
            return self
        hit = set()
        for alias, path, m in self._named():
            for p in m.parameters(recurse=False):
                p.requires_grad_(False)
            if alias in names or path in names:
                for p in m.parameters():
                    p.requires_grad_(False)
                hit.add(alias if alias in names else path)
                if len(hit) == len(names):
                    break
        if len(hit) != len(names):
            raise KeyError(f"no layer named {sorted(set(names) - hit)}")
        return self

    def unfreeze(self, names=None):
        for alias, path, m in self._named():
            if names is None or alias in names or path in names:
                for p in m.parameters():
                    p.requires_grad_(True)
        return self

    def to_keras(self):
        return self

    @property
    def layers(self):
        return [m for _, _, m in self._named() if not list(m.children())]

    def flattened_layers(self, include_container=False):
        return [m for _, _, m in self._named() if include_container or not list(m.children())]


class _Stop(Exception):
    pass


class _Truncated(nn.Module):
    def __init__(self, model, stops):
        super().__init__()
        self.model, self.stops = model, stops

    def forward(self, *xs):
        got, hooks = {}, []
        last = self.stops[-1]

        def mk(i, is_last):
            def hook(_m, _inp, out):
                got[i] = out
                if is_last and len(got) == len(self.stops):
                    raise _Stop()
            return hook

        for i, m in enumerate(self.stops):
            hooks.append(m.register_forward_hook(mk(i, m is last)))
        try:
            self.model(*xs)
        except _Stop:
            pass
        finally:
            for h in hooks:
                h.remove()
        outs = [got[i] for i in range(len(self.stops))]
        return outs[0] if len(outs) == 1 else tuple(outs)
