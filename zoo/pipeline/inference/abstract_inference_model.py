"""The POJO-style serving API of the reference: `AbstractInferenceModel`, `JTensor`, `InferenceModelFactory`, `FloatModel`.

Reference: zoo/src/main/java/com/intel/analytics/zoo/pipeline/inference/AbstractInferenceModel.java (the class web services
extend: `loadBigDL / loadCaffe / loadTensorflow / loadPyTorch / loadOpenVINO`, `reload`, `release`,
`predict(List<Float> input, int... shape)`, `predict(List<List<JTensor>> inputs)`), JTensor.java (data + shape),
InferenceModelFactory.scala / ModelLoader.scala (loaders returning a `FloatModel`), FloatModel.scala:31-88 (model holder with
`predict`, `copy`, `release`), InferenceRuntimeException.java, DeviceType.java.

Everything rides on `InferenceModel` (slot pool + CUDA-graph replay); this file is the calling convention: batches as lists of
records, a record as a list of tensors, a tensor as `JTensor(data, shape)`.
"""
from __future__ import annotations

import copy
from enum import Enum

import numpy as np
import torch

from ...common.utils import JTensor
from .inference_model import InferenceModel


class InferenceRuntimeException(RuntimeError):
    pass


class DeviceType(Enum):
    CPU = "CPU"
    GPU = "GPU"


class AbstractInferenceModel(InferenceModel):
    def __init__(self, concurrent_num=1, auto_scaling_enabled=False, **kw):
        if isinstance(concurrent_num, bool):                     # AbstractInferenceModel(autoScalingEnabled, concurrentNum)
            concurrent_num, auto_scaling_enabled = (auto_scaling_enabled if not isinstance(auto_scaling_enabled, bool) else 1), concurrent_num
        super().__init__(supported_concurrent_num=max(1, int(concurrent_num)), **kw)
        self.auto_scaling_enabled = bool(auto_scaling_enabled)
        self._source = None

    # ---- loaders under their Java names -----------------------------------------------------------------------------------------
    def loadBigDL(self, model_path, weight_path=None):
        self._source = ("load", (model_path, weight_path))
        return self.load(model_path, weight_path)

    def loadCaffe(self, model_path, weight_path=None):
        self._source = ("load_caffe", (model_path, weight_path))
        return self.load_caffe(model_path, weight_path)

    def loadTensorflow(self, model_path, model_type="frozenModel", inputs=None, outputs=None, intra_op_parallelism_threads=1,
                       inter_op_parallelism_threads=1, use_per_session_threads=True):
        self._source = ("load_tensorflow", (model_path, model_type))
        return self.load_tensorflow(model_path, model_type, inputs=inputs, outputs=outputs)

    loadTF = loadTensorflow

    def loadPyTorch(self, model_path_or_bytes):
        import io

        src = io.BytesIO(model_path_or_bytes) if isinstance(model_path_or_bytes, (bytes, bytearray)) else model_path_or_bytes
        try:
            model = torch.jit.load(src, map_location="cpu")
        except Exception:
            if hasattr(src, "seek"):
                src.seek(0)
            model = torch.load(src, map_location="cpu", weights_only=False)
        self._source = ("loadPyTorch", (model_path_or_bytes,))
        return self.load_torch(model)

    def loadOpenVINO(self, model_file_path, weight_file_path, batch_size=0):
        return self.load_openvino(model_file_path, weight_file_path, batch_size)

    def reload(self, model_path, weight_path=None):
        """Swap in a new model file without dropping the serving object."""
        self.release()
        return self.loadBigDL(model_path, weight_path)

    # ---- predict ------------------------------------------------------------------------------------------------------------------
    def predict(self, inputs, *shape):
        """Three calling forms of the reference:
        * `predict(list of floats, d0, d1, ...)` -> list of floats (one tensor in, the first output flattened);
        * `predict([[JTensor, ...], ...])`: a list of records, each a list of input tensors -> a list of records of output tensors;
        * arrays / tensors go to `InferenceModel.predict` unchanged."""
        if self.model is None:
            raise InferenceRuntimeException("no model is loaded")
        if shape:
            x = np.asarray(inputs, dtype=np.float32).reshape([int(s) for s in shape])
            out = super().predict(x)
            return np.asarray(out[0] if isinstance(out, list) else out, dtype=np.float32).reshape(-1).tolist()
        if isinstance(inputs, (list, tuple)) and inputs and isinstance(inputs[0], (list, tuple)) and inputs[0] and isinstance(inputs[0][0], JTensor):
            try:
                n_in = len(inputs[0])
                cols = [np.stack([rec[k].to_ndarray() for rec in inputs]) for k in range(n_in)]      # records -> one batch per input
                out = super().predict(cols if n_in > 1 else cols[0])
            except Exception as e:      # the reference wraps every failure of doPredict
                raise InferenceRuntimeException(f"predict failed: {e}") from e
            outs = out if isinstance(out, list) else [out]
            return [[JTensor.from_ndarray(np.asarray(o[i], dtype=np.float32)) for o in outs] for i in range(len(inputs))]
        return super().predict(inputs)

    def __str__(self):
        return f"AbstractInferenceModel(concurrentNum={self.concurrent_num}, autoScalingEnabled={self.auto_scaling_enabled}, model={type(self.model).__name__})"


class FloatModel:
    """A loaded model + its metadata: what the factory's loaders return and what the model queue of the reference holds."""

    def __init__(self, model, meta_model=None, is_original=True):
        self.model, self.meta_model, self.is_original = model.eval(), meta_model, is_original

    @torch.no_grad()
    def predict(self, inputs):
        if isinstance(inputs, (list, tuple)) and inputs and isinstance(inputs[0], (list, tuple)):
            n_in = len(inputs[0])
            cols = [torch.from_numpy(np.stack([rec[k].to_ndarray() for rec in inputs])) for k in range(n_in)]
            out = self.model(*cols)
            outs = list(out) if isinstance(out, (list, tuple)) else [out]
            return [[JTensor.from_ndarray(o[i].float().cpu().numpy()) for o in outs] for i in range(len(inputs))]
        x = inputs if torch.is_tensor(inputs) else torch.as_tensor(np.asarray(inputs))
        return self.model(x)

    def copy(self, num):
        """`num` clones sharing the weights (parameters are the same tensors; module objects are separate)."""
        out = []
        for _ in range(int(num)):
            memo = {id(p): p for p in list(self.model.parameters()) + list(self.model.buffers())}
            out.append(FloatModel(copy.deepcopy(self.model, memo), self.meta_model, False))
        return out

    def release(self):
        self.model = None

    def is_released(self):
        return self.model is None

    isReleased = is_released

    def __str__(self):
        return f"FloatModel({type(self.model).__name__ if self.model is not None else None}, isOriginal={self.is_original})"


class InferenceModelFactory:
    """Loaders that return a `FloatModel` (InferenceModelFactory.scala:23-120 / ModelLoader.scala)."""

    @staticmethod
    def load_float_model_for_bigdl(model_path, weight_path=None, blas=True):
        from ..api.net import Net

        net = Net.load(model_path, weight_path)
        return FloatModel(net, meta_model=type(net).__name__)

    @staticmethod
    def load_float_model_for_caffe(model_path, weight_path, blas=True):
        from ...models.caffe import load_caffe

        return FloatModel(load_caffe(model_path, weight_path), meta_model="caffe")

    @staticmethod
    def load_float_model_for_tf(model_path, model_type="frozenModel", inputs=None, outputs=None, **_ignored):
        from ..api.net import TFNet

        kind = (model_type or "frozenModel").replace("_", "").lower()
        net = TFNet.from_saved_model(model_path, inputs=inputs, outputs=outputs) if kind == "savedmodel" else TFNet(model_path, inputs, outputs)
        return FloatModel(net, meta_model="tensorflow")

    @staticmethod
    def load_float_model_for_pytorch(model_path_or_bytes):
        m = AbstractInferenceModel()
        m.loadPyTorch(model_path_or_bytes)
        return FloatModel(m.model, meta_model="pytorch")

    loadFloatModelForBigDL, loadFloatModelForCaffe = load_float_model_for_bigdl, load_float_model_for_caffe
    loadFloatModelForTF, loadFloatModelForPyTorch = load_float_model_for_tf, load_float_model_for_pytorch
