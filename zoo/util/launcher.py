"""Local worker launcher: one process per GPU (or per CPU worker) from a single driver process.

This is what `SparkRunner` / `spark-submit` do for the reference (pyzoo/zoo/util/spark.py:27-220: start
executors, ship the environment) reduced to one box: `launch(fn, nprocs)` spawns `nprocs` Python
processes with RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* set, each calls `fn(*args)` (cloudpickled, so
closures and lambdas work like Spark closures do) after joining the process group through
`init_nncontext()`, and the driver gets every rank's return value.  Children die with the driver
(`ProcessMonitor`, the reference's JVMGuard: pyzoo/zoo/ray/raycontext.py:29-48).
"""
from __future__ import annotations

import os
import socket
import traceback

import cloudpickle
import torch.multiprocessing as mp


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, payload, force_cpu, queue):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    if force_cpu:
        os.environ["ZOO_FORCE_CPU"] = "1"
    try:
        fn, args, kwargs = cloudpickle.loads(payload)
        from ..common.nncontext import init_nncontext, stop_orca_context

        init_nncontext()
        out = fn(*args, **kwargs)
        stop_orca_context()
        queue.put((rank, True, cloudpickle.dumps(out)))
    except BaseException as e:  # noqa: BLE001 - report everything to the driver
        queue.put((rank, False, f"{e!r}\n{traceback.format_exc()}"))


def launch(fn, nprocs, args=(), kwargs=None, force_cpu=None, timeout=None):
    """Run `fn(*args, **kwargs)` on `nprocs` ranks; returns the list of results ordered by rank."""
    import torch

    if force_cpu is None:
        force_cpu = not torch.cuda.is_available() or torch.cuda.device_count() < nprocs
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    payload = cloudpickle.dumps((fn, tuple(args), dict(kwargs or {})))
    port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, nprocs, port, payload, force_cpu, q), daemon=True) for r in range(nprocs)]
    for p in procs:
        p.start()
    results, errors = {}, []
    try:
        for _ in range(nprocs):
            rank, ok, blob = q.get(timeout=timeout)
            if ok:
                results[rank] = cloudpickle.loads(blob)
            else:
                errors.append((rank, blob))
                break
    finally:
        for p in procs:
            p.join(timeout=5 if not errors else 0.5)
            if p.is_alive():
                p.terminate()
    if errors:
        raise RuntimeError(f"worker {errors[0][0]} failed:\n{errors[0][1]}")
    return [results[r] for r in range(nprocs)]
