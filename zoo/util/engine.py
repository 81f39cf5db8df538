"""Environment preparation (reference: pyzoo/zoo/util/engine.py:23-170: locate SPARK_HOME / the analytics-zoo jar, put pyspark
and the jar on the path, check Spark versions).  There is no JVM / Spark under zoo-b200: the functions keep their names and
answer for this runtime -- the "classpath" is the native library directory, `prepare_env` makes sure the CUDA library is
built or loadable."""
from __future__ import annotations

import importlib.util
import os


def exist_pyspark():
    return importlib.util.find_spec("pyspark") is not None


def check_spark_source_conflict(spark_home, pyspark_path):
    """The reference warns when the pip-installed pyspark and SPARK_HOME disagree.  Nothing uses Spark here."""
    if spark_home and pyspark_path and not os.path.abspath(pyspark_path).startswith(os.path.abspath(spark_home)):
        import warnings

        warnings.warn(f"SPARK_HOME ({spark_home}) and pyspark ({pyspark_path}) differ; zoo-b200 uses neither")


def get_analytics_zoo_classpath():
    """Directory of the native libraries (libzoo_cuda.so / libzoo_rt.so) -- the analogue of the jar path."""
    from ..native import build

    return str(build.NATIVE_DIR) if hasattr(build, "NATIVE_DIR") else os.path.join(os.path.dirname(os.path.dirname(__file__)), "native")


def compare_version(version1, version2):
    """-1 / 0 / 1 like the reference (dotted numeric versions)."""
    def norm(v):
        return [int(x) for x in str(v).split(".") if x.isdigit()]

    a, b = norm(version1), norm(version2)
    n = max(len(a), len(b))
    a, b = a + [0] * (n - len(a)), b + [0] * (n - len(b))
    return (a > b) - (a < b)


def is_spark_below_2_2():
    if not exist_pyspark():
        return False
    import pyspark

    return compare_version(getattr(pyspark, "__version__", "2.4.3"), "2.2.0") < 0


def prepare_env():
    """Make the runtime importable: the native libraries are located (and reported missing with the build command)."""
    from ..native import lib

    try:
        lib.rt_lib() if hasattr(lib, "rt_lib") else None
    except Exception as e:  # pragma: no cover - depends on the build state
        import warnings

        warnings.warn(f"zoo native runtime not loadable ({e}); run `python -m zoo.native.build`")
