"""TFNet (reference: pyzoo/zoo/pipeline/api/net/ TFNet, zoo/.../pipeline/api/net/TFNet.scala:56-716): runs a frozen
TensorFlow graph through libtensorflow JNI.  There is no TensorFlow runtime on this box (SURVEY 2.3 N2:
dropped as an engine); the class exists so imports resolve and fails with the supported alternative."""
from __future__ import annotations


class TFNet:
    def __init__(self, *a, **kw):
        raise NotImplementedError("TFNet needs the TensorFlow runtime, which is not part of zoo-b200; export the graph to "
                                  "TorchScript/ONNX-free torch code and load it with Net.load_torch or InferenceModel.load_torch")

    from_export_folder = from_session = from_saved_model = staticmethod(lambda *a, **kw: TFNet())
