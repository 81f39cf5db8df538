from .abstract_inference_model import (AbstractInferenceModel, DeviceType, FloatModel, InferenceModelFactory,  # noqa: F401
                                       InferenceRuntimeException)
from .inference_model import InferenceModel  # noqa: F401
from ...common.utils import JTensor  # noqa: F401,E402  (the POJO tensor: data + shape)
