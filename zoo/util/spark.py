"""SparkRunner (reference: pyzoo/zoo/util/spark.py:27-200: create a local or YARN-client SparkContext, conda-pack the
environment, assemble jars).  Workers of zoo-b200 are the GPUs of the box, one process each (`zoo.util.launcher`,
torchrun); the class keeps its two entry points and maps them onto that."""
from __future__ import annotations

from ..common import nncontext


class SparkRunner:
    def __init__(self, spark_log_level="WARN", redirect_spark_log=True):
        self.spark_log_level, self.redirect_spark_log = spark_log_level, redirect_spark_log

    def init_spark_on_local(self, cores, conf=None, python_location=None):
        return nncontext.init_spark_on_local(cores, conf, python_location, self.spark_log_level, self.redirect_spark_log)

    def init_spark_on_yarn(self, hadoop_conf=None, conda_name=None, num_executor=1, executor_cores=1, executor_memory="2g", **kw):
        return nncontext.init_spark_on_yarn(hadoop_conf, conda_name, num_executor, executor_cores, executor_memory, **kw)

    def pack_penv(self, conda_name=None, output=None):
        """Pack what a worker on another machine needs to run this framework (reference: `conda pack` of the named environment into
        `<tmp>/python_env.tar.gz`, spark.py:100-113, shipped to the YARN executors).  Workers here are processes of the same image
        (docker/hyperzoo on Kubernetes), so the archive holds the `zoo` package with its built native libraries, the launch
        scripts and a lock file of the installed distributions (`requirements.lock`, from importlib.metadata) to reproduce the
        Python environment with pip; `conda_name`, when given and conda-pack is installed, packs that conda environment instead."""
        import importlib.metadata as md
        import io
        import os
        import shutil
        import subprocess
        import tarfile
        import tempfile

        out = output or os.path.join(tempfile.mkdtemp(), "python_env.tar.gz")
        if conda_name and shutil.which("conda-pack"):
            subprocess.check_call(["conda-pack", "--format", "tar.gz", "-n", conda_name, "-f", "-o", out])
            return out
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        skip = ("__pycache__", ".pytest_cache", ".git")
        with tarfile.open(out, "w:gz") as tar:
            for top in ("zoo", "scripts", "include", "setup.py"):
                path = os.path.join(root, top)
                if os.path.exists(path):
                    tar.add(path, arcname=top, filter=lambda ti: None if any(part in skip for part in ti.name.split("/")) else ti)
            lock = "\n".join(sorted(f"{d.metadata['Name']}=={d.version}" for d in md.distributions() if d.metadata["Name"])) + "\n"
            info = tarfile.TarInfo("requirements.lock")
            data = lock.encode()
            info.size = len(data)
            tar.addfile(info, io.BytesIO(data))
        return out
