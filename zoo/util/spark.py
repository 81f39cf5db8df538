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

    def pack_penv(self, conda_name):
        raise NotImplementedError("conda-pack of the environment is a YARN deployment step; zoo-b200 runs one process per GPU of one box "
                                  "(use scripts/zoo-submit or torchrun)")
