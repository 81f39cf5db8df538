from ..unsupported import GANEstimator  # noqa: F401
