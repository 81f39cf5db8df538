from .image_config import ImageConfigure, PaddingParam  # noqa: F401
from .image_model import ImageModel  # noqa: F401
