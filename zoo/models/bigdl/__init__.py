"""BigDL model files (`Module.saveModule` / `ZooModel.saveModel`) <-> torch modules.  See `bigdl_loader` and `bigdl_writer`."""
from .bigdl_loader import BigDLConversionException, BigDLGraph, BigDLLoader, load_bigdl, register_bigdl_layer  # noqa: F401
from .bigdl_writer import save_bigdl  # noqa: E402,F401
