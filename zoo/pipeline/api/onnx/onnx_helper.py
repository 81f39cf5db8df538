"""Attribute / tensor helpers (reference: pyzoo/zoo/pipeline/api/onnx/onnx_helper.py)."""
from .onnx_loader import _attrs, _tensor_to_numpy


class OnnxHelper:
    @staticmethod
    def parse_attr(attr_proto):
        return _attrs(type("N", (), {"attribute": attr_proto})())

    @staticmethod
    def to_numpy(tensor_proto):
        return _tensor_to_numpy(tensor_proto)

    @staticmethod
    def get_shape_from_node(valueInfoProto):
        return [int(d.dim_value) for d in valueInfoProto.type.tensor_type.shape.dim]
