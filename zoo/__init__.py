"""zoo-b200: a Blackwell-native framework with the capabilities of Analytics Zoo.

Public package layout follows the reference's Python package (``pyzoo/zoo``) so user code
keeps its imports (``zoo.common.nncontext``, ``zoo.pipeline.estimator``,
``zoo.models.recommendation`` ...); the engine underneath is PyTorch + hand-written sm_100a
CUDA kernels (``csrc/``) + NVLink peer-memory collectives.
"""
__version__ = "0.8.0.dev0+b200"

from .common.nncontext import (  # noqa: F401,E402
    init_nncontext, init_spark_on_local, init_spark_on_yarn, init_orca_context, stop_orca_context, get_context,
)
