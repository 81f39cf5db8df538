"""Per-operator mappers (reference: pyzoo/zoo/pipeline/api/onnx/mapper/*.py, one `XxxMapper(OperatorMapper)` per ONNX
operator that builds a zoo keras layer).  Here every mapper is a thin handle on the interpreter's implementation of
that operator (`OnnxModel._op_Xxx`), kept so code that looks mappers up by name (`OperatorMapper.of(node)`) works."""
from __future__ import annotations

import numpy as np
import torch

from ..onnx_loader import OnnxModel, _attrs

OPERATORS = ["Abs", "Add", "AveragePool", "BatchNormalization", "Cast", "Clip", "Concat", "Constant", "Conv", "Div", "Dropout", "Elu", "Exp",
             "Flatten", "Gather", "Gemm", "GlobalAveragePool", "Greater", "HardSigmoid", "LeakyRelu", "Log", "LogSoftmax", "LRN", "MatMul", "MaxPool",
             "Mul", "Neg", "Pow", "ReduceMean", "ReduceSum", "Relu", "Reshape", "Shape", "Sigmoid", "Slice", "Softmax", "Sqrt", "Squeeze", "Sub",
             "Tanh", "Transpose", "Unsqueeze"]


class OperatorMapper:
    op_type = None

    def __init__(self, node, initializer=None, inputs=None):
        self.node, self.initializer, self.inputs = node, initializer or {}, inputs or []
        self.onnx_attr = _attrs(node)

    @staticmethod
    def of(node, initializer=None, inputs=None):
        cls = globals().get(node.op_type + "Mapper")
        if cls is None:
            raise NotImplementedError(f"no mapper for ONNX operator {node.op_type}")
        return cls(node, initializer, inputs)

    def run(self, *tensors):
        m = OnnxModel.__new__(OnnxModel)
        torch.nn.Module.__init__(m)
        res = getattr(m, "_op_" + self.op_type)([torch.as_tensor(np.asarray(t)) if not torch.is_tensor(t) else t for t in tensors], self.onnx_attr)
        return res

    to_tensor = run


for _op in OPERATORS:
    globals()[_op + "Mapper"] = type(_op + "Mapper", (OperatorMapper,), {"op_type": _op})
del _op
