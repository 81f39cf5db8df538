"""Runtime bring-up: `init_nncontext` and friends on a single multi-GPU box.

The reference creates a SparkContext with zoo's conf, exports KMP/OMP settings and calls BigDL
`Engine.init` (pyzoo/zoo/common/nncontext.py:23-222, zoo/.../common/NNContext.scala:133-236).
Here the "cluster" is the GPUs of one box: the context discovers them, reads the launcher's
RANK / LOCAL_RANK / WORLD_SIZE, binds this process to its GPU and (when world_size > 1) joins a
``torch.distributed`` process group -- NCCL on GPUs (bootstrap + baseline collectives), gloo on
CPU (plumbing tests).  The names and the call shapes of the reference are kept.
"""
from __future__ import annotations

import logging
import os
from dataclasses import dataclass, field

import torch

log = logging.getLogger("zoo")
_CONTEXT = None

ZOO_CONF_DEFAULTS = {
    # reference: zoo/src/main/resources/spark-analytics-zoo.conf + bigdl.* JVM properties
    "zoo.failure.retryTimes": 5,
    "zoo.failure.retryTimeInterval": 120,
    "zoo.versionCheck": False,
    "zoo.comm.backend": "auto",        # auto | p2p | nccl | gloo
    "zoo.comm.bucket_mb": 25,
}


@dataclass
class ZooContext:
    rank: int = 0
    local_rank: int = 0
    world_size: int = 1
    device: torch.device = field(default_factory=lambda: torch.device("cpu"))
    backend: str | None = None
    conf: dict = field(default_factory=dict)
    cores: int = 1
    app_name: str = "zoo-b200"
    owns_process_group: bool = False

    @property
    def is_distributed(self):
        return self.world_size > 1

    @property
    def is_gpu(self):
        return self.device.type == "cuda"

    def barrier(self):
        if self.is_distributed:
            import torch.distributed as dist

            dist.barrier()

    def getConf(self):
        return self

    def get(self, key, default=None):
        return self.conf.get(key, default)

    def stop(self):
        stop_orca_context()


def init_env(cores=None):
    """Thread settings for the host side (reference: init_env sets KMP_AFFINITY / OMP_NUM_THREADS /
    ZOO_NUM_MKLTHREADS, nncontext.py:162-195).  The env knobs keep their names."""
    n = os.environ.get("ZOO_NUM_MKLTHREADS") or os.environ.get("OMP_NUM_THREADS")
    if n and n != "all":
        try:
            torch.set_num_threads(max(1, int(n)))
        except (ValueError, RuntimeError):
            pass
    elif cores:
        try:
            torch.set_num_threads(max(1, int(cores)))
        except RuntimeError:
            pass


def init_spark_conf(conf=None):
    merged = dict(ZOO_CONF_DEFAULTS)
    if conf:
        merged.update(conf if isinstance(conf, dict) else dict(conf))
    for k in list(merged):
        env = "ZOO_" + k.split("zoo.", 1)[-1].replace(".", "_").upper()
        if env in os.environ:
            merged[k] = type(merged[k])(os.environ[env]) if not isinstance(merged[k], bool) else os.environ[env] == "1"
    return merged


def check_version():
    """Reference compares Spark/BigDL versions (nncontext.py:225-260); here: CUDA arch sanity."""
    if torch.cuda.is_available():
        major, minor = torch.cuda.get_device_capability()
        if (major, minor) != (10, 0):
            log.warning("zoo-b200 kernels are built for sm_100a; found sm_%d%d", major, minor)


def init_nncontext(conf=None, app_name="zoo-b200", cluster_mode=None, cores=None, device=None, backend=None, **_ignored):
    """Create (or return) the process-wide context.

    conf: dict of zoo.* options or a string app name (the reference accepts either).
    """
    global _CONTEXT
    if isinstance(conf, str):
        conf, app_name = None, conf
    if _CONTEXT is not None:
        return _CONTEXT
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if device is None:
        use_gpu = torch.cuda.is_available() and os.environ.get("ZOO_FORCE_CPU") != "1"
        device = torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1)) if use_gpu else torch.device("cpu")
    else:
        device = torch.device(device)
    if device.type == "cuda":
        torch.cuda.set_device(device)
    owns = False
    if world > 1:
        import torch.distributed as dist

        if not dist.is_initialized():
            backend = backend or ("nccl" if device.type == "cuda" else "gloo")
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            kw = {"device_id": device} if backend == "nccl" else {}
            dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
            owns = True
        else:
            backend = dist.get_backend()
    init_env(cores)
    check_version()
    _CONTEXT = ZooContext(rank, local_rank, world, device, backend, init_spark_conf(conf), cores or os.cpu_count() or 1,
                          app_name, owns)
    return _CONTEXT


def get_context():
    return _CONTEXT if _CONTEXT is not None else init_nncontext()


def init_spark_on_local(cores=2, conf=None, python_location=None, spark_log_level="WARN", redirect_spark_log=True):
    """Reference: local[cores] SparkContext (nncontext.py:23-39).  Here: the local box."""
    return init_nncontext(conf, cores=None if cores == "*" else cores)


def init_spark_on_yarn(hadoop_conf=None, conda_name=None, num_executor=1, executor_cores=1, executor_memory="2g", **kw):
    """YARN launch (nncontext.py:42-101) has no meaning on one 8-GPU box: workers are the GPUs.
    Kept so scripts run unchanged; `num_executor` is only checked against the GPU count."""
    ctx = init_nncontext(kw.get("conf"))
    if torch.cuda.is_available() and num_executor > torch.cuda.device_count():
        log.warning("num_executor=%d > %d GPUs on this box", num_executor, torch.cuda.device_count())
    return ctx


def init_orca_context(cluster_mode="local", cores=2, memory="2g", num_nodes=1, init_ray_on_spark=False, **kw):
    return init_nncontext(kw.get("conf"), cores=None if cores == "*" else cores)


def stop_orca_context():
    global _CONTEXT
    if _CONTEXT is not None and _CONTEXT.owns_process_group:
        import torch.distributed as dist

        if dist.is_initialized():
            dist.destroy_process_group()
    _CONTEXT = None


def getOrCreateSparkContext(conf=None, appName=None):
    return init_nncontext(conf, appName or "zoo-b200")


def get_analytics_zoo_conf():
    """The framework's default configuration (reference: reads spark-analytics-zoo.conf from the jar)."""
    return dict(ZOO_CONF_DEFAULTS)


def load_conf(conf_str):
    """`key value` lines -> dict (nncontext.py `load_conf`)."""
    out = {}
    for line in conf_str.split("\n"):
        line = line.strip()
        if line and not line.startswith("#"):
            k, _, v = line.partition(" ")
            out[k] = v.strip()
    return out
