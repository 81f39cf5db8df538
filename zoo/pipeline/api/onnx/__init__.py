from .onnx_loader import OnnxLoader, OnnxModel  # noqa: F401
