from .inference_model import InferenceModel  # noqa: F401
