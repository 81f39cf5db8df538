"""Rank watchdog: heartbeats in shared memory, a timeout on steps that hang, and an abort that lets the retry loop restart from
the newest checkpoint.

The reference gets failure detection from Spark (task re-execution) and adds a retry-from-checkpoint loop
(Topology.scala:1180-1262: `bigdl.failure.retryTimes` failures within `retryTimes x retryTimeInterval` seconds); Ray processes are
reaped by `ProcessMonitor` (pyzoo/zoo/ray/process.py:139-150).  One process per GPU has no scheduler above it, so a rank that dies
or hangs inside a collective must be noticed by its peers: every rank bumps a counter in a small memory-mapped file each step; a
daemon thread on every rank checks all counters and, when one has not moved for `timeout` seconds (or the own step has been
running that long), runs `on_failure` -- by default: log, then abort the process so the launcher / retry loop takes over
(device-side spins are bounded on their own: the mbarrier and flag waits in the kernels trap after a clock budget).
"""
from __future__ import annotations

import logging
import mmap
import os
import struct
import tempfile
import threading
import time

log = logging.getLogger("zoo.watchdog")
_SLOT = 32       # bytes per rank: u64 step counter, f64 wall time of the last beat, u64 pid, u64 state (0 run, 1 done)


class HeartbeatBoard:
    """`world` slots in one memory-mapped file (on /dev/shm when present): lock-free, one writer per slot."""

    def __init__(self, world, rank, path=None, create=None):
        self.world, self.rank = int(world), int(rank)
        if path is None:
            base = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
            tag = os.environ.get("MASTER_PORT", "0") + "-" + os.environ.get("TORCHELASTIC_RUN_ID", str(os.getppid()))
            path = os.path.join(base, f"zoo-heartbeat-{tag}")
        self.path = path
        size = _SLOT * self.world
        create = (self.rank == 0) if create is None else create
        if create or not os.path.exists(path):
            with open(path, "ab") as fh:
                if fh.tell() < size:
                    fh.write(b"\0" * (size - fh.tell()))
        self._fh = open(path, "r+b")
        self._mm = mmap.mmap(self._fh.fileno(), size)
        self.beat(0)

    def beat(self, step, state=0):
        struct.pack_into("<QdQQ", self._mm, self.rank * _SLOT, int(step), time.time(), os.getpid(), int(state))

    def read(self, rank):
        step, t, pid, state = struct.unpack_from("<QdQQ", self._mm, rank * _SLOT)
        return {"step": step, "time": t, "pid": pid, "state": state}

    def snapshot(self):
        return [self.read(r) for r in range(self.world)]

    def close(self, unlink=False):
        try:
            self.beat(self.read(self.rank)["step"], state=1)
            self._mm.close()
            self._fh.close()
            if unlink and os.path.exists(self.path):
                os.unlink(self.path)
        except (ValueError, OSError):
            pass


def _pid_alive(pid):
    if pid <= 0:
        return True
    try:
        os.kill(pid, 0)
        return True
    except ProcessLookupError:
        return False
    except PermissionError:
        return True


class RankWatchdog:
    """Start with `RankWatchdog(world, rank, timeout=...).start()`, call `beat(step)` once per step, `stop()` at the end.

    A peer is declared failed when its process is gone, or when it is still marked running and has not beaten for `timeout`
    seconds while this rank has.  The own rank is declared hung when ITS last beat is older than `timeout` (a step that never
    returns, e.g. a collective waiting for a dead peer)."""

    def __init__(self, world, rank, timeout=300.0, interval=None, on_failure=None, path=None, board=None):
        self.board = board or HeartbeatBoard(world, rank, path)
        self.world, self.rank, self.timeout = int(world), int(rank), float(timeout)
        self.interval = interval if interval is not None else max(0.05, min(5.0, self.timeout / 10))
        self.on_failure = on_failure or self._abort
        self.failed = None
        self._stop = threading.Event()
        self._thread = None
        self._armed = False

    # ---- the training loop's side ----------------------------------------------------------------------------------------
    def beat(self, step):
        self._armed = True
        self.board.beat(step)

    def start(self):
        if self._thread is None:
            self._thread = threading.Thread(target=self._run, name="zoo-watchdog", daemon=True)
            self._thread.start()
        return self

    def stop(self, unlink=None):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2 * self.interval + 1)
            self._thread = None
        self.board.close(unlink=(self.rank == 0) if unlink is None else unlink)

    # ---- the monitor ------------------------------------------------------------------------------------------------------
    def check(self, now=None):
        """One pass over the board; returns None or a dict describing the failure."""
        now = time.time() if now is None else now
        snap = self.board.snapshot()
        me = snap[self.rank]
        if self._armed and me["state"] == 0 and now - me["time"] > self.timeout:
            return {"kind": "hang", "rank": self.rank, "step": me["step"], "silent_for": now - me["time"]}
        for r, s in enumerate(snap):
            if r == self.rank or s["pid"] == 0 or s["state"] == 1:
                continue
            if not _pid_alive(s["pid"]):
                return {"kind": "dead", "rank": r, "step": s["step"], "pid": s["pid"]}
            if self._armed and now - s["time"] > self.timeout and now - me["time"] <= self.timeout:
                return {"kind": "stalled", "rank": r, "step": s["step"], "silent_for": now - s["time"]}
        return None

    def _run(self):
        while not self._stop.wait(self.interval):
            f = self.check()
            if f is not None:
                self.failed = f
                try:
                    self.on_failure(f)
                finally:
                    return

    def _abort(self, failure):
        log.error("watchdog on rank %d: %s -- aborting so the job can restart from the newest checkpoint", self.rank, failure)
        os._exit(75)      # EX_TEMPFAIL: the launcher's retry loop (zoo.failure.retryTimes) decides what happens next
