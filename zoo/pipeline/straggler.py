"""Gradient dropping: straggler mitigation of the synchronous step.

Reference: the zoo optimizer hands `dropPercentage`, `warmupIterationNum`, `computeThresholdbatchSize` and `maxDropPercentage`
to the distributed step (zoo/src/main/scala/com/intel/analytics/zoo/pipeline/api/keras/models/Topology.scala:1096-1100).  There
every node runs several model replicas as tasks on slices of the mini-batch; after a warm-up a time threshold is taken from the
recorded task times so that `dropPercentage` of the tasks would exceed it, tasks still running at the threshold are abandoned,
the gradient is divided by the number of replicas that finished, and an iteration in which more than `maxDropPercentage` of the
replicas were abandoned is discarded.

Here a rank is one GPU and its unit of droppable work is a MICRO-BATCH: the local batch is split into `micro_batches` slices whose
gradients accumulate in the flat gradient buffer.  The step has a deadline `micro_batches x threshold`; a rank starts its next
micro-batch only if, at its observed pace, that micro-batch finishes before the deadline (the first one always runs).  Ranks then
exchange (micro-batches done, samples done, seconds per micro-batch) in one small all-gather, the accumulated gradient SUM is
divided by the samples that were really processed on all ranks, and the exchange + optimizer step runs as usual -- or the
iteration is discarded on every rank when too much was dropped.  The threshold is re-estimated every `batch_size` iterations from
the gathered times, the way the reference does (k-th largest time with k = dropPercentage x window, less what was already
dropped in the window; grown by 1 % when the window dropped enough).

Cost: one device synchronisation per micro-batch (the pace has to be observed on the host) and no CUDA-graph replay of the step,
so this is off unless `TrainingEngine.set_drop_module_property` / conf `zoo.drop.percentage` turns it on.
"""
import logging
import time

import torch

log = logging.getLogger("zoo.straggler")


def kth_largest(values, k):
    """k-th largest element, k counted from 1 (k <= 0 gives the largest, k > len the smallest)."""
    s = sorted(values, reverse=True)
    return s[min(max(k, 1), len(s)) - 1]


class GradientDropper:
    def __init__(self, ctx, drop_percentage, max_drop_percentage, batch_size=100, warmup_iteration=200, micro_batches=4,
                 clock=time.perf_counter):
        if not 0.0 <= drop_percentage < 1.0 or not 0.0 <= max_drop_percentage < 1.0:
            raise ValueError("drop_percentage and max_drop_percentage are fractions in [0, 1)")
        if max_drop_percentage < drop_percentage:
            raise ValueError("max_drop_percentage must be >= drop_percentage (an iteration that drops what was asked for must still update)")
        if micro_batches < 2:
            raise ValueError("gradient dropping needs at least 2 micro-batches per rank (one always runs, the others may be dropped)")
        self.ctx = ctx
        self.drop_percentage, self.max_drop_percentage = float(drop_percentage), float(max_drop_percentage)
        self.window, self.warmup, self.micro_batches = int(batch_size), int(warmup_iteration), int(micro_batches)
        self.clock = clock
        self.delay = None            # test hook: callable(rank, micro_batch) -> seconds of extra work
        self.threshold = None        # seconds per micro-batch; None until the first window after warm-up is complete
        self.iteration = 0           # steps seen by this object
        self._times = []             # seconds per micro-batch, one entry per (rank, iteration) of the current window
        self._dropped_in_window = 0  # (rank, iteration) entries of the current window that dropped at least one micro-batch
        self.stats = {"dropped_micro_batches": 0, "discarded_iterations": 0, "iterations": 0, "threshold_updates": 0}

    # -- inside one step ---------------------------------------------------------------------------------------------------
    def begin(self):
        self.iteration += 1
        self._t0 = self.clock()
        self._done = 0

    def may_start(self, j):
        """True when micro-batch `j` (0-based) of this step should run."""
        if j == 0 or self.threshold is None:
            return True
        elapsed = self.clock() - self._t0
        pace = elapsed / j
        return elapsed + pace <= self.micro_batches * self.threshold

    def finished(self, j):
        if self.delay is not None:
            d = self.delay(self.ctx.rank, j)
            if d:
                time.sleep(d)
        self._done = j + 1

    def exchange(self, samples_done):
        """All ranks: returns (samples processed on all ranks, True when the iteration should update the weights)."""
        m, W = self.micro_batches, self.ctx.world_size
        per_mb = (self.clock() - self._t0) / max(self._done, 1)
        mine = torch.tensor([float(self._done), float(samples_done), per_mb], dtype=torch.float64)
        if W > 1:
            import torch.distributed as dist

            dev = self.ctx.device if dist.get_backend() == "nccl" else torch.device("cpu")
            rows = [torch.empty(3, dtype=torch.float64, device=dev) for _ in range(W)]
            dist.all_gather(rows, mine.to(dev))
            table = torch.stack(rows).cpu()
        else:
            table = mine.view(1, 3)
        done_all, samples_all = int(table[:, 0].sum().item()), int(table[:, 1].sum().item())
        dropped = W * m - done_all
        self.stats["iterations"] += 1
        self.stats["dropped_micro_batches"] += dropped
        if self.iteration > self.warmup:
            self._times.extend(table[:, 2].tolist())
            self._dropped_in_window += int((table[:, 0] < m).sum().item())
            if len(self._times) >= self.window * W:
                self._update_threshold()
        update = dropped <= self.max_drop_percentage * W * m
        if not update:
            self.stats["discarded_iterations"] += 1
            log.warning("iteration %d: %d of %d micro-batches were dropped (more than maxDropPercentage %.2f); the gradients of this "
                        "iteration are discarded", self.iteration, dropped, W * m, self.max_drop_percentage)
        return samples_all, update

    def _update_threshold(self):
        # the reference's rule with a (rank, iteration) entry as the task: k = dropPercentage x window entries may be slow; what the
        # window already dropped is taken off, and a window that dropped its share only relaxes the threshold by 1 %
        k = int(self.drop_percentage * len(self._times))
        if self.drop_percentage > 0.0:
            if self.threshold is None or k > self._dropped_in_window:
                kk = k - (self._dropped_in_window if self.threshold is not None else 0)
                self.threshold = kth_largest(self._times, kk) if kk >= 1 else max(self._times) * 1.01
            else:
                self.threshold *= 1.01
            self.stats["threshold_updates"] += 1
            log.info("gradient dropping: threshold %.3f ms per micro-batch (window of %d entries, %d of them dropped work)",
                     self.threshold * 1e3, len(self._times), self._dropped_in_window)
        self._times, self._dropped_in_window = [], 0


def split_batch(x, m):
    """Split a tensor / tuple of tensors along dim 0 into at most `m` near-equal slices; returns a list of the same structure."""
    if x is None:
        return [None] * m
    if isinstance(x, (tuple, list)):
        parts = [split_batch(t, m) for t in x]
        return [tuple(p[i] for p in parts) for i in range(len(parts[0]))]
    n = x.shape[0]
    step = -(-n // m)
    return [x[i:i + step] for i in range(0, n, step)]


def batch_rows(x):
    while isinstance(x, (tuple, list)):
        x = x[0]
    return int(x.shape[0])
