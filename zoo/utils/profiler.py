"""Tracing helpers: wall-clock wrappers with the reference's log lines, CUDA-event phase timers and NVTX ranges.

Reference: `Utils.timeIt` (zoo/.../common/Utils.scala:40-47), `InferenceSupportive.timing`
(pipeline/inference/InferenceSupportive.scala:41-48), `EstimateSupportive.throughputing / throughputingWithLoss`
(pipeline/estimator/EstimateSupportive.scala:36-65: "... Throughput is X records/second").  Those are wall-clock only; on a GPU
the phases of a step are timed with CUDA events on the launching stream (no host sync inside the step) and the slowest rank's
time is what is reported.
"""
from __future__ import annotations

import contextlib
import logging
import time
from collections import OrderedDict

import torch

log = logging.getLogger("zoo.profiler")


@contextlib.contextmanager
def time_it(name, logger=None):
    """`timeIt(name){...}`: logs "<name> time elapsed [<s> s, <ms> ms]."."""
    t0 = time.perf_counter()
    try:
        yield
    finally:
        dt = time.perf_counter() - t0
        (logger or log).info("%s time elapsed [%d s, %.3f ms].", name, int(dt), dt * 1e3)


timing = time_it


@contextlib.contextmanager
def throughputing(name, batch_size, logger=None, loss=None):
    """`throughputing(name, batchSize){...}`: "<name> time elapsed [...], throughput: X records/second[. Loss is L]"."""
    t0 = time.perf_counter()
    try:
        yield
    finally:
        dt = max(time.perf_counter() - t0, 1e-12)
        tail = "" if loss is None else f" Loss is {float(loss() if callable(loss) else loss):.6f}."
        (logger or log).info("%s time elapsed [%d s, %.3f ms], throughput: %.1f records/second.%s", name, int(dt), dt * 1e3, batch_size / dt, tail)


@contextlib.contextmanager
def nvtx_range(name):
    """NVTX range when CUDA is available (shows up in nsys / ncu timelines); a no-op otherwise."""
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class PhaseTimer:
    """Per-phase device time of a step without synchronising inside it.

        pt = PhaseTimer(device)
        with pt.phase("forward"): ...
        with pt.phase("backward"): ...
        pt.step()                       # end of one step (events are resolved lazily)
        pt.summary()  ->  {"forward": {"mean_ms": .., "max_ms": .., "count": ..}, ...}

    On a CUDA device every phase is bracketed by two events recorded on the current stream (and an NVTX range); events are
    only read in `summary()`, after one synchronize.  On the CPU it falls back to `perf_counter`.  `summary(reduce=True)` takes
    the MAX over ranks of every mean (the slowest rank bounds a synchronous step)."""

    def __init__(self, device=None):
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.cuda = self.device.type == "cuda"
        self._pending = []              # (name, start_event, end_event)
        self._done = OrderedDict()      # name -> list of ms
        self.steps = 0

    @contextlib.contextmanager
    def phase(self, name):
        if self.cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.nvtx.range_push(name)
            e0.record()
            try:
                yield
            finally:
                e1.record()
                torch.cuda.nvtx.range_pop()
                self._pending.append((name, e0, e1))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self._done.setdefault(name, []).append((time.perf_counter() - t0) * 1e3)

    def step(self):
        self.steps += 1
        if len(self._pending) > 4096:
            self._resolve()

    def _resolve(self):
        if self._pending:
            torch.cuda.synchronize(self.device)
            for name, e0, e1 in self._pending:
                self._done.setdefault(name, []).append(e0.elapsed_time(e1))
            self._pending = []

    def summary(self, reduce=False, skip_first=0):
        self._resolve()
        out = OrderedDict()
        for name, ms in self._done.items():
            ms = ms[skip_first:] or ms
            out[name] = {"mean_ms": sum(ms) / len(ms), "max_ms": max(ms), "count": len(ms)}
        if reduce and torch.distributed.is_available() and torch.distributed.is_initialized():
            t = torch.tensor([v["mean_ms"] for v in out.values()], dtype=torch.float64, device=self.device if self.cuda else "cpu")
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            for v, m in zip(out.values(), t.tolist()):
                v["mean_ms_max_over_ranks"] = m
        return out

    def report(self, logger=None, **kw):
        s = self.summary(**kw)
        tot = sum(v["mean_ms"] for v in s.values()) or 1.0
        for name, v in s.items():
            (logger or log).info("phase %-16s mean %.3f ms  max %.3f ms  (%4.1f %%, %d samples)", name, v["mean_ms"], v["max_ms"],
                                 100 * v["mean_ms"] / tot, v["count"])
        return s

    def reset(self):
        self._pending, self._done, self.steps = [], OrderedDict(), 0
