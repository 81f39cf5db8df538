"""Nested-structure helpers (reference: pyzoo/zoo/util/nest.py:19-80): `flatten`, `pack_sequence_as`,
`ptensor_to_numpy`."""
from __future__ import annotations


def flatten(seq):
    if isinstance(seq, dict):
        out = []
        for k in sorted(seq):
            out += flatten(seq[k])
        return out
    if isinstance(seq, (list, tuple)):
        out = []
        for s in seq:
            out += flatten(s)
        return out
    return [seq]


def pack_sequence_as(structure, flat_sequence):
    flat = list(flat_sequence)
    pos = [0]

    def build(s):
        if isinstance(s, dict):
            return {k: build(s[k]) for k in sorted(s)}
        if isinstance(s, (list, tuple)):
            r = [build(x) for x in s]
            return type(s)(r) if not hasattr(s, "_fields") else type(s)(*r)
        if pos[0] >= len(flat):
            raise ValueError(f"structure has more leaves than the {len(flat)} values given")
        v = flat[pos[0]]
        pos[0] += 1
        return v

    out = build(structure)
    if pos[0] != len(flat):
        raise ValueError(f"structure has {pos[0]} leaves but {len(flat)} values were given")
    return out


def ptensor_to_numpy(seq):
    import torch

    return pack_sequence_as(seq, [t.detach().cpu().numpy() if torch.is_tensor(t) else t for t in flatten(seq)])


def is_sequence(s):
    return isinstance(s, (list, tuple, dict))
