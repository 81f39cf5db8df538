"""`TensorOperation`: broadcasting helpers of the reference's layer-norm internals (reference:
zoo/src/main/scala/com/intel/analytics/zoo/common/TensorOperation.scala:25-110: `expandSize`, `expandTensor`, `subTensor`,
`divTensor`; used by InternalLayerNorm.scala:44-49).  BigDL tensors have no implicit broadcasting, so the reference builds
stride-0 views by hand; torch has the views built in, and these functions keep the reference's contract -- trailing dimensions
are aligned, a dimension of size 1 stretches, the FIRST argument is updated in place and returned."""
from __future__ import annotations

import torch


def expand_size(tensor, other):
    """The broadcast size of the two tensors (trailing dimensions aligned)."""
    try:
        return list(torch.broadcast_shapes(tuple(tensor.shape), tuple(other.shape)))
    except RuntimeError:
        raise ValueError(f"tensor size not match {'x'.join(map(str, tensor.shape))} {'x'.join(map(str, other.shape))}") from None


def expand_tensor(tensor, other):
    """`other` viewed at the broadcast size (no copy); `tensor` is grown in place when the broadcast size is larger than it."""
    size = expand_size(tensor, other)
    if list(tensor.shape) != size:
        tensor.set_(tensor.expand(size).contiguous())
    return other.expand(size)


def sub_tensor(tensor, other):
    return tensor.sub_(expand_tensor(tensor, other))


def div_tensor(tensor, other):
    return tensor.div_(expand_tensor(tensor, other))


class TensorOperation:
    expandSize = expand_size = staticmethod(expand_size)
    expandTensor = expand_tensor = staticmethod(expand_tensor)
    subTensor = sub_tensor = staticmethod(sub_tensor)
    divTensor = div_tensor = staticmethod(div_tensor)
