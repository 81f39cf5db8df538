from .raycontext import RayContext  # noqa: F401
