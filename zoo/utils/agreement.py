"""Step-boundary agreement between the ranks of one box: "does everybody start step n, or did somebody fail?"

The reference's retry loop lives in the Spark driver, which sees a failed task and re-initialises every executor from the
newest checkpoint (Topology.scala:1180-1262).  One process per GPU has no driver, so the ranks agree among themselves: before
every step each rank posts `(sequence, status)` into its slot of a small memory-mapped board and waits until every slot has
reached the same sequence number.  A rank that caught an exception posts FAILED for that sequence instead; all ranks then see
the same verdict for the same step and enter recovery together (reload the newest checkpoint, Engine.train).  The wait is a
handful of shared-memory reads when ranks run in lock step (they do: every step ends in a collective), and it never blocks the
GPU, whose queue the host keeps ahead of.  Failures that strike INSIDE a step (a rank dies in a kernel or a collective) cannot
be agreed on in-process; those end in the watchdog's abort and a restart of the job, which resumes from the newest checkpoint.
"""
from __future__ import annotations

import mmap
import os
import struct
import tempfile
import time

_SLOT = 16          # u64 sequence, u64 status
OK, FAILED = 0, 1


class PeerFailure(RuntimeError):
    """Raised on the healthy ranks when a peer reported a failure at a step boundary."""


class StepAgreement:
    @staticmethod
    def default_path(tag="agree"):
        base = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
        job = os.environ.get("MASTER_PORT", "0") + "-" + os.environ.get("TORCHELASTIC_RUN_ID", str(os.getppid()))
        return os.path.join(base, f"zoo-{tag}-{job}-{os.environ.get('TORCHELASTIC_RESTART_COUNT', '0')}")

    @staticmethod
    def reset(path, world):
        """Create the board with every slot at sequence 0 (one rank does this BEFORE the others open it; a board left over from
        an earlier job with the same name would otherwise let the first waits pass on stale sequence numbers)."""
        with open(path, "wb") as fh:
            fh.write(b"\0" * (_SLOT * int(world)))

    def __init__(self, world, rank, path=None, timeout=120.0, tag="agree"):
        self.world, self.rank, self.timeout = int(world), int(rank), float(timeout)
        if path is None:
            path = self.default_path(tag)
        self.path = path
        size = _SLOT * self.world
        with open(path, "ab") as fh:            # every rank may create it; appends of zeros are idempotent up to `size`
            if fh.tell() < size:
                fh.write(b"\0" * (size - fh.tell()))
        self._fh = open(path, "r+b")
        self._mm = mmap.mmap(self._fh.fileno(), size)
        self.seq = 0

    def _post(self, seq, status):
        # status first, sequence last: a reader that sees the new sequence also sees its status
        struct.pack_into("<Q", self._mm, self.rank * _SLOT + 8, int(status))
        struct.pack_into("<Q", self._mm, self.rank * _SLOT, int(seq))

    def _collect(self, seq):
        """Wait until every rank has posted `seq`; returns the list of ranks that reported FAILED."""
        t0, spins = time.time(), 0
        while True:
            slots = [struct.unpack_from("<QQ", self._mm, r * _SLOT) for r in range(self.world)]
            if all(s >= seq for s, _ in slots):
                return [r for r, (s, st) in enumerate(slots) if s == seq and st == FAILED]
            spins += 1
            if spins > 200:
                time.sleep(0.0005)
            if time.time() - t0 > self.timeout:
                late = [r for r, (s, _) in enumerate(slots) if s < seq]
                raise TimeoutError(f"ranks {late} did not reach step boundary {seq} within {self.timeout:.0f} s")

    def check_in(self):
        """Healthy path: post OK for the next boundary and wait for the peers.  Raises PeerFailure if somebody failed."""
        self.seq += 1
        self._post(self.seq, OK)
        bad = self._collect(self.seq)
        if bad:
            raise PeerFailure(f"rank(s) {bad} failed at step boundary {self.seq}")

    def report_failure(self):
        """Failing path (called instead of `check_in` for this boundary): tell the peers, then wait for them."""
        self.seq += 1
        self._post(self.seq, FAILED)
        self._collect(self.seq)

    def close(self, unlink=False):
        try:
            self._mm.close()
            self._fh.close()
            if unlink and os.path.exists(self.path):
                os.unlink(self.path)
        except (ValueError, OSError):
            pass
