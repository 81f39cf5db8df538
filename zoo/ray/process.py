"""Child-process bookkeeping (reference: pyzoo/zoo/ray/process.py:27-150 `ProcessInfo`, `ProcessMonitor`;
raycontext.py:29-48 `JVMGuard`): every worker PID is registered and killed when the driver exits."""
from __future__ import annotations

import atexit
import os
import signal


class ProcessInfo:
    def __init__(self, out="", err="", errorcode=0, pid=None, tag="default", pgid=None, node_ip="127.0.0.1"):
        self.out, self.err, self.errorcode, self.pid, self.tag, self.pgid, self.node_ip = out, err, errorcode, pid, tag, pgid, node_ip
        self.master_addr = None

    def __str__(self):
        return f"node_ip: {self.node_ip} tag: {self.tag}, pgid: {self.pgid}, pid: {self.pid}, returncode: {self.errorcode}"


class ProcessMonitor:
    _registered = set()
    _hooked = False

    @classmethod
    def register_pids(cls, pids):
        cls._registered.update(int(p) for p in pids)
        if not cls._hooked:
            atexit.register(cls.clean_fn)
            cls._hooked = True

    @classmethod
    def unregister(cls, pids):
        cls._registered.difference_update(int(p) for p in pids)

    @classmethod
    def clean_fn(cls):
        for pid in list(cls._registered):
            try:
                os.kill(pid, signal.SIGTERM)
            except (ProcessLookupError, PermissionError):
                pass
        cls._registered.clear()


JVMGuard = ProcessMonitor
