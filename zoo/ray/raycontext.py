"""RayContext: the worker pool of the box.

Reference: pyzoo/zoo/ray/raycontext.py:190-337 starts `ray start --head` / raylets inside Spark barrier
tasks (one per executor) and `ray.init`s the driver; everything above it (XShards, AutoML trials, the
MXNet trainer) only needs three things from Ray: run a function on a worker (`remote`), keep state in
a worker (`actor`), and fetch results (`get`).  Ray is not installed here, and on one box the workers
are local processes: `RayContext(...).init()` starts a process pool (spawned, cloudpickled functions,
one optional GPU per worker), registered with `ProcessMonitor` so they die with the driver.
`local_mode=True` (or `num_workers=0`) runs everything inline, which is what the reference's
`RayContext(local_ray_node_num=...)` local mode is used for in tests.
"""
from __future__ import annotations

import itertools
import os
import threading
from concurrent.futures import Future, ProcessPoolExecutor

import cloudpickle
import multiprocessing as mp

from .process import ProcessMonitor

_ACTORS = {}     # worker-side registry: actor id -> instance


def _init_worker(gpu_ids, counter):
    with counter.get_lock():
        slot = counter.value
        counter.value += 1
    if gpu_ids:
        os.environ["CUDA_VISIBLE_DEVICES"] = str(gpu_ids[slot % len(gpu_ids)])
    os.environ["ZOO_WORKER_SLOT"] = str(slot)


def _call(payload):
    fn, args, kwargs = cloudpickle.loads(payload)
    return cloudpickle.dumps(fn(*args, **kwargs))


def _actor_new(aid, payload):
    cls, args, kwargs = cloudpickle.loads(payload)
    _ACTORS[aid] = cls(*args, **kwargs)
    return cloudpickle.dumps(True)


def _actor_call(aid, name, payload):
    args, kwargs = cloudpickle.loads(payload)
    return cloudpickle.dumps(getattr(_ACTORS[aid], name)(*args, **kwargs))


class ObjectRef:
    def __init__(self, fut, decode=True):
        self._fut, self._decode = fut, decode

    def get(self, timeout=None):
        r = self._fut.result(timeout)
        return cloudpickle.loads(r) if self._decode else r


class ActorHandle:
    """Methods return ObjectRefs; the instance lives in ONE worker process (a single-process executor)."""

    def __init__(self, ctx, cls, args, kwargs):
        self._ctx = ctx
        self._id = next(ctx._ids)
        self._local = None
        if ctx.local_mode:
            self._local = cls(*args, **kwargs)
        else:
            self._exec = ProcessPoolExecutor(1, mp_context=mp.get_context("spawn"))
            self._exec.submit(_actor_new, self._id, cloudpickle.dumps((cls, args, kwargs))).result()
            ProcessMonitor.register_pids(list(self._exec._processes.keys()))
            ctx._actors.append(self)

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)

        def remote(*args, **kwargs):
            if self._local is not None:
                f = Future()
                try:
                    f.set_result(getattr(self._local, name)(*args, **kwargs))
                except Exception as e:  # noqa: BLE001
                    f.set_exception(e)
                return ObjectRef(f, decode=False)
            return ObjectRef(self._exec.submit(_actor_call, self._id, name, cloudpickle.dumps((args, kwargs))))

        remote.remote = remote
        return remote

    def kill(self):
        if self._local is None and self._exec is not None:
            pids = list(self._exec._processes.keys())
            self._exec.shutdown(wait=True, cancel_futures=True)
            ProcessMonitor.unregister(pids)
            self._exec = None


class RayContext:
    _active = None

    def __init__(self, sc=None, redis_port=None, password="123456", object_store_memory=None, verbose=False, env=None,
                 extra_params=None, local_ray_node_num=None, num_workers=None, gpus=None, local_mode=False):
        self.sc = sc
        self.num_workers = num_workers if num_workers is not None else (local_ray_node_num or min(4, os.cpu_count() or 1))
        self.local_mode = local_mode or self.num_workers == 0
        self.env = dict(env or {})
        self.gpus = list(gpus) if gpus is not None else []
        self.initialized = False
        self._pool = None
        self._ids = itertools.count()
        self._lock = threading.Lock()
        self._actors = []

    @classmethod
    def get(cls, initialize=True):
        if cls._active is None:
            raise Exception("No active RayContext. Please create a RayContext and init it first")
        if initialize and not cls._active.initialized:
            cls._active.init()
        return cls._active

    def init(self, driver_cores=0):
        if self.initialized:
            return self
        os.environ.update({k: str(v) for k, v in self.env.items()})
        if not self.local_mode:
            ctx = mp.get_context("spawn")
            counter = ctx.Value("i", 0)
            self._pool = ProcessPoolExecutor(self.num_workers, mp_context=ctx, initializer=_init_worker, initargs=(self.gpus, counter))
            list(self._pool.map(_call, [cloudpickle.dumps((os.getpid, (), {}))] * self.num_workers))   # start the workers now
            ProcessMonitor.register_pids(list(self._pool._processes.keys()))
        self.initialized = True
        RayContext._active = self
        return self

    # -- the three things the framework needs from "Ray" ---------------------------------------
    def remote(self, fn, *args, **kwargs) -> ObjectRef:
        if not self.initialized:
            self.init()
        if self.local_mode:
            f = Future()
            try:
                f.set_result(fn(*args, **kwargs))
            except Exception as e:  # noqa: BLE001
                f.set_exception(e)
            return ObjectRef(f, decode=False)
        return ObjectRef(self._pool.submit(_call, cloudpickle.dumps((fn, args, kwargs))))

    def map(self, fn, items):
        return [r.get() for r in [self.remote(fn, it) for it in items]]

    def actor(self, cls, *args, **kwargs) -> ActorHandle:
        if not self.initialized:
            self.init()
        return ActorHandle(self, cls, args, kwargs)

    @staticmethod
    def get_objects(refs, timeout=None):
        return [r.get(timeout) for r in refs] if isinstance(refs, (list, tuple)) else refs.get(timeout)

    def stop(self):
        for a in self._actors:          # actor processes end with the context (the reference kills raylets and their actors)
            a.kill()
        self._actors = []
        if self._pool is not None:
            pids = list(self._pool._processes.keys())
            self._pool.shutdown(wait=True, cancel_futures=True)
            ProcessMonitor.unregister(pids)
            self._pool = None
        self.initialized = False
        if RayContext._active is self:
            RayContext._active = None

    def purge(self):
        self.stop()


JVMGuard = ProcessMonitor


class RayServiceFuncGenerator:
    """In the reference: builds the `ray start` shell commands run inside Spark barrier tasks (raycontext.py:51-187).
    Workers here are local processes started by `RayContext.init()`, so this only records the settings."""

    def __init__(self, python_loc=None, redis_port=None, ray_node_cpu_cores=None, password=None, object_store_memory=None,
                 verbose=False, env=None, extra_params=None):
        self.python_loc, self.redis_port, self.ray_node_cpu_cores = python_loc, redis_port, ray_node_cpu_cores
        self.password, self.object_store_memory, self.verbose, self.env, self.extra_params = password, object_store_memory, verbose, env, extra_params
