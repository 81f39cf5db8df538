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


def pids_from_gpid(gpid):
    """PIDs whose process group is `gpid` (reads /proc)."""
    out = []
    for d in os.listdir("/proc"):
        if d.isdigit():
            try:
                if os.getpgid(int(d)) == int(gpid):
                    out.append(int(d))
            except (ProcessLookupError, PermissionError):
                pass
    return out


def session_execute(command, env=None, tag=None, fail_fast=False, timeout=120):
    """Run a shell command in its own session; returns a ProcessInfo (process.py:60-88)."""
    import subprocess

    p = subprocess.Popen(command, shell=True, env=env, cwd=None, stdout=subprocess.PIPE, stderr=subprocess.PIPE, start_new_session=True)
    try:
        out, err = p.communicate(timeout=timeout)
    except subprocess.TimeoutExpired:
        p.kill()
        out, err = p.communicate()
    info = ProcessInfo(out.decode(errors="replace"), err.decode(errors="replace"), p.returncode, p.pid, tag or "default", os.getpgid(os.getpid()))
    if fail_fast and p.returncode != 0:
        raise Exception(f"command failed ({p.returncode}): {command}\n{info.err}")
    return info
