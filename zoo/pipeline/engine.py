"""The training / evaluation / prediction engine behind Estimator, KerasNet and NNEstimator.

This is the B200 counterpart of `InternalDistriOptimizer` + BigDL `DistriOptimizer`
(zoo/src/main/scala/.../pipeline/api/keras/models/Topology.scala:963-1600) and `LocalEstimator`
(pipeline/estimator/LocalEstimator.scala).  Behaviour kept from the reference:

* one engine object per model; calling `train` again continues epochs / iterations / optimizer
  state (Estimator.scala:124-131, DistriEstimatorSpec "Train multi times");
* end / checkpoint / validation are driven by ZooTriggers over the state table
  (`epoch`, `neval`, `Loss`, `score`); checkpoint and validation share one trigger
  (Topology.scala:1329-1342);
* checkpoint layout ``<modelDir>/<yyyy-MM-dd_HH-mm-ss>/models/model.<neval>`` +
  ``optimMethod-<name>.<neval>``, summaries under ``<modelDir>/<ts>/summary/{train,validation}``
  (Topology.scala:1294-1306,1170-1176);
* retry-from-last-checkpoint on failure: at most `retryTimes` (5) failures inside a
  `retryTimes x retryTimeInterval` (120 s) window, `ValueError`/`TypeError` are never retried
  (Topology.scala:1180-1262);
* per-iteration log line with throughput in records/second (EstimateSupportive.scala:36-65) and
  TensorBoard scalars `Loss`, `Throughput`, `LearningRate`.

What is different: there are no Spark jobs.  One process drives one GPU; the step is
forward -> fused loss -> backward (tcgen05 GEMMs, sparse embedding updates inside backward) ->
gradient exchange fused with the optimizer (`zoo.parallel.dp`) -> next batch, whose host-to-device
copy was started a step earlier on a side stream.
"""
from __future__ import annotations

import glob
import logging
import os
import re
import time
from datetime import datetime

import numpy as np
import torch
import torch.nn as nn

from ..common.nncontext import get_context
from ..common.triggers import EveryEpoch, MaxEpoch, ZooTrigger
from ..feature.common import FeatureSet, MiniBatch
from ..native import lib as _nl
from ..parallel.flat import FusedOptimizer, OptConfig
from ..tensorboard import TrainSummary, ValidationSummary
from .api.keras.metrics import to_metric
from .api.keras.objectives import to_loss
from .api.keras.optimizers import to_opt_config

log = logging.getLogger("zoo.engine")


def _to_device(x, device, non_blocking=True):
    if x is None:
        return None
    if isinstance(x, (tuple, list)):
        return tuple(_to_device(t, device, non_blocking) for t in x)
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    return x.to(device, non_blocking=non_blocking) if torch.is_tensor(x) else x


def _nbytes(x):
    if x is None:
        return 0
    if isinstance(x, (tuple, list)):
        return sum(_nbytes(t) for t in x)
    return x.numel() * x.element_size() if torch.is_tensor(x) and not x.is_cuda else 0


class Prefetcher:
    """Overlap the H2D copy of batch i+1 with the compute of batch i (pinned source -> async DMA
    on a side stream).  Device-resident FeatureSets pass straight through."""

    _streams = {}       # one copy stream per device, shared by the passes of a run (creating a stream per pass costs ~1 ms)

    def __init__(self, it, device):
        self.it, self.device = iter(it), device
        self.stream = None
        if device.type == "cuda":
            key = device.index if device.index is not None else torch.cuda.current_device()
            self.stream = Prefetcher._streams.get(key)
            if self.stream is None:
                self.stream = Prefetcher._streams[key] = torch.cuda.Stream(device)
        self.h2d_bytes = 0
        self._next = None
        self._preload()

    def _preload(self):
        try:
            b = next(self.it)
        except StopIteration:
            self._next = None
            return
        if self.stream is None:
            self._next = MiniBatch(_to_device(b.inputs, self.device), _to_device(b.targets, self.device), b.size)
            return
        self.h2d_bytes = _nbytes(b.inputs) + _nbytes(b.targets)
        with torch.cuda.stream(self.stream):
            self._next = MiniBatch(_to_device(b.inputs, self.device), _to_device(b.targets, self.device), b.size)
            if self.h2d_bytes:
                b._h2d_done = torch.cuda.Event()      # lets the producer of `b` recycle its pinned staging buffer
                b._h2d_done.record(self.stream)

    def __iter__(self):
        return self

    def __next__(self):
        if self._next is None:
            raise StopIteration
        if self.stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
            for t in _flatten(self._next.inputs) + _flatten(self._next.targets):
                if torch.is_tensor(t) and t.is_cuda:
                    t.record_stream(torch.cuda.current_stream(self.device))
        b = self._next
        self._preload()
        return b


def _flatten(x):
    if x is None:
        return []
    if isinstance(x, (tuple, list)):
        out = []
        for t in x:
            out += _flatten(t)
        return out
    return [x]


class TrainingEngine:
    """Owns model placement, the flat-parameter arena, the fused optimizer, triggers, summaries
    and checkpoints for one model."""

    def __init__(self, model: nn.Module, optim_methods=None, model_dir=None, ctx=None, grad_sync="auto"):
        self.ctx = ctx or get_context()
        self.model = model
        self.model_dir = model_dir
        self.optim_methods = optim_methods
        self.grad_sync = grad_sync
        self.fopt: FusedOptimizer | None = None
        self.state = {"epoch": 1, "neval": 1, "Loss": float("inf"), "score": float("-inf"), "recordsProcessedThisEpoch": 0}
        self.train_summary = None
        self.val_summary = None
        self.run_dir = None
        self.ckpt_dir = None
        self.overwrite_checkpoint = True
        self.clip = None
        self._dp = None
        self._tb_dir = None
        self.metrics_history = []
        self.last_throughput = 0.0
        self.on_step = None          # optional callback(iteration, loss_tensor) after every step
        self.async_loss = False      # read the loss back with a one-step lag (no pipeline stall)
        self.watchdog = None         # RankWatchdog when ZOO_WATCHDOG_TIMEOUT / zoo.watchdog.timeout is set
        self.use_cuda_graph = None   # None = auto (models that declare `graph_safe = True`)
        self._graph = None
        self._loss_ring = None
        self._last_criterion = None
        self.autocast_dtype = None   # e.g. torch.bfloat16: run generic torch models' forward under autocast (masters stay fp32)
        self.channels_last = False
        self._dropper = None         # GradientDropper when set_drop_module_property / conf zoo.drop.percentage turned it on

    # ----------------------------------------------------------------- set-up --
    @property
    def device(self):
        return self.ctx.device

    def _default_cfg_and_map(self):
        om = self.optim_methods
        if om is None:
            return OptConfig("sgd", lr=1e-2, b1=0.0), {}
        if isinstance(om, dict):
            m = {k: to_opt_config(v) for k, v in om.items()}
            names = {n.split(".")[0] for n, _ in self.model.named_parameters()}
            default = next(iter(m.values()))
            # a single entry keyed by the model's own name (reference: Map(model.getName -> optim))
            if len(m) == 1 and not (set(m) & names):
                return default, {}
            return default, m
        return to_opt_config(om), {}

    def _prepare_tables_for_dp(self):
        """Replicated sparse tables cannot run their private optimizer when ranks see different
        batches: train them as dense parameters so their gradients are reduced with the rest."""
        from ..ops.embedding import EmbeddingTable

        if self.ctx.world_size > 1:
            for m in self.model.modules():
                if isinstance(m, EmbeddingTable) and m.sparse and m.world == 1:
                    m.sparse = False
                    m.weight.requires_grad_(True)

    def prepare(self):
        if self.fopt is not None:
            return self
        self.model.to(self.device)
        if self.channels_last and self.device.type == "cuda":
            self.model.to(memory_format=torch.channels_last)
        self._prepare_tables_for_dp()
        self._sync_sparse_optimizers()
        default, per = self._default_cfg_and_map()
        from ..parallel.dp import make_grad_sync

        self._dp = make_grad_sync(self.ctx, self.grad_sync)
        self.fopt = FusedOptimizer.build(self.model, default, per, device=self.device, allocator=self._dp.allocator(),
                                         pad_to=self._dp.pad_to())
        self._dp.attach(self.fopt)
        self._alias_shadows()
        if self.clip is not None:
            self._apply_clip()
        if self.ctx.world_size > 1:
            self._dp.broadcast_parameters()
            import torch.distributed as dist

            for b in self.model.buffers():        # e.g. BatchNorm statistics start identical on every rank
                if b.is_floating_point() or b.dtype in (torch.int64, torch.int32):
                    dist.broadcast(b, src=0)
        return self

    def _sync_sparse_optimizers(self):
        """Sparse embedding tables take the engine's default optimizer family and lr."""
        from ..ops.embedding import EmbeddingTable

        default, _ = self._default_cfg_and_map()
        fam = {"adam": "adam", "adam_zoo": "adam", "adamw": "adam", "adamw_bert": "adam", "adagrad": "adagrad"}.get(default.name, "sgd")
        for m in self.model.modules():
            if isinstance(m, EmbeddingTable) and m.fused_sparse and not getattr(m, "_user_opt", False):
                m.optimizer = fam
                m.opt_cfg.update(lr=default.lr, betas=(default.b1, default.b2) if fam == "adam" else m.opt_cfg["betas"],
                                 eps=default.eps if fam != "sgd" else m.opt_cfg["eps"], weight_decay=default.weight_decay)
                m._state_ready = False

    def _alias_shadows(self):
        from ..ops.dense import FusedDense

        a = self.fopt.arena
        self._dense_layers = []
        for m in self.model.modules():
            if isinstance(m, FusedDense):
                sh = a.shadow_of(m.weight) if m.weight.requires_grad else None
                aliased = sh is not None and m.in_features % 8 == 0
                if aliased:
                    m.weight_bf16 = sh
                    m.mark_shadow_fresh()
                self._dense_layers.append((m, aliased))

    def _after_step_shadows(self):
        for m, aliased in self._dense_layers:
            if aliased:
                m.mark_shadow_fresh()
            else:
                m.refresh_shadow(force=True)

    # --------------------------------------------------------------- clipping --
    def set_constant_gradient_clipping(self, min, max):
        self.clip = ("const", float(min), float(max))
        self._apply_clip()

    def set_gradient_clipping_by_l2_norm(self, clip_norm):
        self.clip = ("l2", float(clip_norm))
        self._apply_clip()

    def clear_gradient_clipping(self):
        self.clip = None
        if self.fopt is not None:
            self.fopt.clear_gradient_clipping()

    def _apply_clip(self):
        if self.fopt is None or self.clip is None:
            return
        if self.clip[0] == "const":
            self.fopt.set_constant_gradient_clipping(self.clip[1], self.clip[2])
        else:
            self.fopt.set_gradient_clipping_by_l2_norm(self.clip[1])

    # ------------------------------------------------------- summaries / ckpt --
    def set_tensorboard(self, log_dir, app_name):
        self._tb_dir = (log_dir, app_name)
        if self.ctx.rank == 0:
            self.train_summary = TrainSummary(log_dir, app_name)
            self.val_summary = ValidationSummary(log_dir, app_name)

    def set_checkpoint(self, path, over_write=True):
        self.ckpt_dir, self.overwrite_checkpoint = path, over_write
        os.makedirs(path, exist_ok=True)

    def _ensure_run_dir(self):
        if self.model_dir is None or self.run_dir is not None:
            return
        ts = datetime.now().strftime("%Y-%m-%d_%H-%M-%S")
        if self.ctx.world_size > 1:      # one run directory per job: rank 0's timestamp
            import torch.distributed as dist

            box = [ts]
            dist.broadcast_object_list(box, src=0)
            ts = box[0]
        self.run_dir = os.path.join(self.model_dir, ts)
        if self.ckpt_dir is None:
            self.set_checkpoint(os.path.join(self.run_dir, "models"))
        if self.train_summary is None and self.ctx.rank == 0:
            self.train_summary = TrainSummary(self.run_dir, "summary")
            self.val_summary = ValidationSummary(self.run_dir, "summary")

    def _ckpt_paths(self, neval):
        gname = lambda g: "optimMethod-" + (type(self.model).__name__ if g.name == "None" else g.name)
        return os.path.join(self.ckpt_dir, f"model.{neval}"), [os.path.join(self.ckpt_dir, f"{gname(g)}.{neval}") for g in self.fopt.groups]

    def _sharded_tables(self):
        from ..ops.embedding import EmbeddingTable

        return [(n, m) for n, m in self.model.named_modules() if isinstance(m, EmbeddingTable) and m.world > 1]

    def save_checkpoint(self):
        """Write `model.<neval>` + `optimMethod-<name>.<neval>` (rank 0) and, when tables are row-sharded, one
        `model.<neval>.shard-<r>-of-<W>` per rank with that rank's rows and their optimizer state.  Collective when
        world_size > 1: every rank must call it (the owner-computes optimizer state is gathered first)."""
        if self.ckpt_dir is None:
            return None
        ctx, W = self.ctx, self.ctx.world_size
        neval = self.state["neval"] - 1
        mpath, opaths = self._ckpt_paths(neval)
        sharded = self._sharded_tables()
        if W > 1 and hasattr(self._dp, "gather_optimizer_state"):
            self._dp.gather_optimizer_state()       # K1 keeps m / v of a range only on its owner: make rank 0's copy complete
        if ctx.rank == 0 and self.overwrite_checkpoint:
            for f in glob.glob(os.path.join(self.ckpt_dir, "model.*")) + glob.glob(os.path.join(self.ckpt_dir, "optimMethod-*")):
                os.remove(f)
        if W > 1:
            ctx.barrier()                           # old files are gone before any rank writes its shard
        skip = {f"{n}.weight" if n else "weight" for n, _ in sharded}
        if sharded:
            blob = {}
            for n, m in sharded:
                m._ensure_state()
                blob[n] = {"weight": m.weight.detach().cpu(), "rows_local": m.rows_local, "num_embeddings": m.num_embeddings,
                           "exp_avg": None if m.exp_avg is None else m.exp_avg.cpu(),
                           "exp_avg_sq": None if m.exp_avg_sq is None else m.exp_avg_sq.cpu(), "hyper": m.hyper.cpu()}
            torch.save({"world": W, "rank": ctx.rank, "tables": blob}, f"{mpath}.shard-{ctx.rank}-of-{W}")
        if ctx.rank == 0:
            torch.save({"model": {k: v.detach().cpu() for k, v in self.model.state_dict().items() if k not in skip},
                        "sparse": self._sparse_state(), "state": dict(self.state), "world": W,
                        "sharded_tables": [n for n, _ in sharded]}, mpath)
            osd = self.fopt.state_dict()
            for g, p in zip(self.fopt.groups, opaths):
                torch.save({"group": g.name, "range": (g.start, g.end), "optim": osd}, p)
        if W > 1:
            ctx.barrier()                           # the checkpoint is complete on disk before anyone trains on
        return mpath

    def _sparse_state(self):
        from ..ops.embedding import EmbeddingTable

        out = {}
        for n, m in self.model.named_modules():
            if isinstance(m, EmbeddingTable) and m.world == 1 and m.fused_sparse and m._state_ready:
                out[n] = {"exp_avg": None if m.exp_avg is None else m.exp_avg.cpu(),
                          "exp_avg_sq": None if m.exp_avg_sq is None else m.exp_avg_sq.cpu(), "hyper": m.hyper.cpu()}
        return out

    @staticmethod
    def _model_files(d):
        return [f for f in glob.glob(os.path.join(d, "model.*")) if re.fullmatch(r"model\.\d+", os.path.basename(f))]

    def latest_checkpoint(self):
        if self.ckpt_dir is None:
            return None
        files = self._model_files(self.ckpt_dir)
        return max(files, key=lambda f: (os.path.getmtime(f), int(f.rsplit(".", 1)[-1]))) if files else None

    def find_resume_checkpoint(self):
        """Newest `model.<n>` under any earlier run directory of `model_dir` (used when a restarted job resumes)."""
        if self.model_dir is None or not os.path.isdir(self.model_dir):
            return None
        files = []
        for d in glob.glob(os.path.join(self.model_dir, "*", "models")):
            files += self._model_files(d)
        return max(files, key=lambda f: (os.path.getmtime(f), int(f.rsplit(".", 1)[-1]))) if files else None

    def load_checkpoint(self, path=None):
        """Restore model, optimizer state and trigger state.  Collective when world_size > 1: every rank loads the replicated
        part from `model.<n>` and its OWN row shard from `model.<n>.shard-<rank>-of-<world>`."""
        path = path or self.latest_checkpoint()
        if path is None:
            raise FileNotFoundError("no checkpoint to load")
        self.prepare()
        ctx, W = self.ctx, self.ctx.world_size
        blob = torch.load(path, map_location="cpu", weights_only=False)
        sharded = self._sharded_tables()
        saved_sharded = blob.get("sharded_tables", [])
        if sharded or saved_sharded:
            if blob.get("world", 1) != W or sorted(saved_sharded) != sorted(n for n, _ in sharded):
                raise ValueError(f"checkpoint {path} was written with world_size={blob.get('world', 1)} and sharded tables "
                                 f"{saved_sharded}; this job has world_size={W} and {[n for n, _ in sharded]} (re-sharding a "
                                 "checkpoint across a different number of GPUs is not supported)")
        skip = {f"{n}.weight" if n else "weight" for n, _ in sharded}
        with torch.no_grad():
            own = self.model.state_dict()
            for k, v in blob["model"].items():
                if k in own and k not in skip:
                    own[k].copy_(v)
        neval = path.rsplit(".", 1)[-1]
        ofiles = sorted(glob.glob(os.path.join(os.path.dirname(path), f"optimMethod-*.{neval}")))
        if ofiles:
            self.fopt.load_state_dict(torch.load(ofiles[0], map_location="cpu", weights_only=False)["optim"])
        from ..ops.embedding import EmbeddingTable

        if sharded:
            sp = f"{path}.shard-{ctx.rank}-of-{W}"
            if not os.path.exists(sp):
                raise FileNotFoundError(f"row-shard file {sp} of the checkpoint is missing")
            sb = torch.load(sp, map_location="cpu", weights_only=False)["tables"]
            with torch.no_grad():
                for n, m in sharded:
                    st = sb[n]
                    if st["rows_local"] != m.rows_local or st["num_embeddings"] != m.num_embeddings:
                        raise ValueError(f"shard of table {n} has {st['rows_local']} rows, expected {m.rows_local}")
                    m._ensure_state()
                    m.weight.data.copy_(st["weight"])
                    if st["exp_avg"] is not None and m.exp_avg is not None:
                        m.exp_avg.copy_(st["exp_avg"])
                    if st["exp_avg_sq"] is not None and m.exp_avg_sq is not None:
                        m.exp_avg_sq.copy_(st["exp_avg_sq"])
                    m.hyper.copy_(st["hyper"])
            for n, m in sharded:
                m.refresh_shadow()                  # collective: peers never gather a half-restored shard
        for n, m in self.model.named_modules():
            if isinstance(m, EmbeddingTable) and m.world == 1 and n in blob.get("sparse", {}):
                m._ensure_state()
                st = blob["sparse"][n]
                if st["exp_avg"] is not None:
                    m.exp_avg.copy_(st["exp_avg"])
                if st["exp_avg_sq"] is not None:
                    m.exp_avg_sq.copy_(st["exp_avg_sq"])
                m.hyper.copy_(st["hyper"])
                m.refresh_shadow()
        self.state.update(blob.get("state", {}))
        if self.fopt.arena.flat_p16 is not None:
            self.fopt.arena.flat_p16.copy_(self.fopt.arena.flat_p)
        self._after_step_shadows()
        if W > 1:
            if self.device.type == "cuda":
                torch.cuda.synchronize()
            if hasattr(self._dp, "after_restore"):
                self._dp.after_restore()
            ctx.barrier()
        return self

    # --------------------------------------------------------------- the step --
    def forward_loss(self, criterion, x, y):
        """Fused path: model exposes logits and the criterion owns a logits kernel."""
        m = self.model
        kind = m.logits_kind() if hasattr(m, "logits_kind") else None
        if kind is not None and kind == getattr(criterion, "logits_kind", None):
            if hasattr(m, "training_loss") and torch.is_grad_enabled() and not isinstance(x, tuple):
                return m.training_loss(x, y, criterion), None      # the model may fuse its head with the loss (zoo.ops.head_loss)
            logits = m.forward_logits(*x) if isinstance(x, tuple) else m.forward_logits(x)
            return criterion.from_logits(logits, y), logits
        if self.autocast_dtype is not None and self.device.type == "cuda":
            if self.channels_last and torch.is_tensor(x) and x.dim() == 4:
                x = x.contiguous(memory_format=torch.channels_last)
            with torch.autocast("cuda", dtype=self.autocast_dtype):
                out = m(*x) if isinstance(x, tuple) else m(x)
            out = out.float() if torch.is_tensor(out) else out
            return criterion(out, y), out
        out = m(*x) if isinstance(x, tuple) else m(x)
        return criterion(out, y), out

    def _eager_step(self, criterion, inputs, targets):
        from ..ops import head_loss as _head_loss
        from ..ops import loss as _fused_loss

        # forward + backward of THIS step: gradients were zeroed by the previous optimizer step and backward is seeded with 1, so
        # a fused head may accumulate its weight gradients while it computes the loss
        _head_loss.ENGINE_STEP["on"] = os.environ.get("ZOO_FUSED_HEAD", "1") != "0"
        if hasattr(self._dp, "begin_step") and os.environ.get("ZOO_HYPER_PREFETCH", "1") != "0":
            self._dp.begin_step()      # hyper-parameter refresh on a side stream, next to forward / backward
        try:
            loss, _ = self.forward_loss(criterion, inputs, targets)
            reg = self.model.regularization_loss() if hasattr(self.model, "regularization_loss") else None
            if reg is not None:
                loss = loss + reg
            _fused_loss.UNIT_UPSTREAM["on"] = True   # we seed backward with exactly 1.0
            loss.backward()
        finally:
            _fused_loss.UNIT_UPSTREAM["on"] = False
            _head_loss.ENGINE_STEP["on"] = False
        self._dp.step()          # gradient exchange fused with the optimizer
        self._after_step_shadows()
        return loss.detach()

    # ------------------------------------------------------- gradient dropping --
    def set_drop_module_property(self, drop_percentage, max_drop_percentage, batch_size=100, warmup_iteration=200, micro_batches=4):
        """Straggler mitigation (reference: dropPercentage / maxDropPercentage / computeThresholdbatchSize / warmupIterationNum,
        Topology.scala:1096-1100).  See `zoo.pipeline.straggler`.  `drop_percentage` 0 turns it off."""
        from .straggler import GradientDropper

        self._dropper = None if not drop_percentage else GradientDropper(self.ctx, drop_percentage, max_drop_percentage, batch_size,
                                                                         warmup_iteration, micro_batches)
        return self

    def _drop_step(self, criterion, inputs, targets):
        """The step with droppable micro-batches: gradient SUMS accumulate in the flat buffer, ranks agree on what was processed,
        the sum is divided by the samples that were really processed, then the usual exchange + optimizer step."""
        from ..ops.embedding import EmbeddingTable
        from .straggler import batch_rows, split_batch

        d = self._dropper
        if any(isinstance(m, EmbeddingTable) and m.sparse for m in self.model.modules()):
            raise ValueError("gradient dropping works on the dense gradient exchange; tables with their own sparse optimizer update inside "
                             "backward and cannot take back a dropped micro-batch (build the model with sparse=False or turn dropping off)")
        xs, ys = split_batch(inputs, d.micro_batches), split_batch(targets, d.micro_batches)
        cuda = self.device.type == "cuda"
        d.begin()
        samples, loss_sum = 0, None
        for j, (x, y) in enumerate(zip(xs, ys)):
            if not d.may_start(j):
                break
            n = batch_rows(x)
            loss, _ = self.forward_loss(criterion, x, y)
            (loss * n).backward()            # criterions average over their batch: x n gives the sum this micro-batch contributes
            loss_sum = loss.detach() * n if loss_sum is None else loss_sum + loss.detach() * n
            samples += n
            if cuda:
                torch.cuda.synchronize()     # the pace is observed on the host
            d.finished(j)
        samples_all, update = d.exchange(samples)
        reg = self.model.regularization_loss() if hasattr(self.model, "regularization_loss") else None
        if not update:
            self.fopt.zero_grad()
            return loss_sum / max(samples, 1)
        self.fopt.arena.flat_g.mul_(self.ctx.world_size / max(samples_all, 1))   # the exchange averages over ranks
        if reg is not None:
            reg.backward()
        self._dp.step()
        self._after_step_shadows()
        return loss_sum / max(samples, 1)

    def _graph_enabled(self):
        if self.device.type != "cuda" or self._dropper is not None:
            return False
        if self.use_cuda_graph is not None:
            return bool(self.use_cuda_graph)
        return bool(getattr(self.model, "graph_safe", False)) and self._dp.graph_safe()

    def train_step(self, criterion, batch: MiniBatch):
        """One optimizer step.  Launch-bound models replay the whole step (forward, loss, backward,
        sparse + dense optimizers) from ONE CUDA graph: ~40 launches become one `cudaGraphLaunch`."""
        if self._dropper is not None:
            return self._drop_step(criterion, batch.inputs, batch.targets)
        if not self._graph_enabled():
            return self._eager_step(criterion, batch.inputs, batch.targets)
        g = self._graph
        key = (id(criterion), tuple((tuple(t.shape), t.dtype) for t in _flatten(batch.inputs) + _flatten(batch.targets)))
        if g is None or g["key"] != key:
            g = self._graph = self._capture(criterion, batch, key)
            if g["count"] <= g["warm"]:
                return g["loss"]
        if g["graph"] is None:          # still warming up eagerly
            g["count"] += 1
            loss = self._eager_step(criterion, batch.inputs, batch.targets)
            if g["count"] > g["warm"]:
                self._finish_capture(criterion, batch)
            return loss
        for dst, src in zip(g["static"], _flatten(batch.inputs) + _flatten(batch.targets)):
            dst.copy_(src, non_blocking=True)
        self.fopt.pre_graph_replay()
        g["graph"].replay()
        _nl.LAUNCH_COUNT["n"] += g["launches"]   # kernels of ours inside the replayed graph
        return g["loss"]

    def _capture(self, criterion, batch, key):
        g = {"key": key, "graph": None, "count": 1, "warm": 3, "static": None, "loss": None, "launches": 0}
        g["loss"] = self._eager_step(criterion, batch.inputs, batch.targets)
        return g

    def _finish_capture(self, criterion, batch):
        g = self._graph
        clone = lambda x: None if x is None else (tuple(t.clone() for t in x) if isinstance(x, (tuple, list)) else x.clone())
        s_in, s_tg = clone(batch.inputs), clone(batch.targets)
        g["static"] = _flatten(s_in) + _flatten(s_tg)
        torch.cuda.synchronize()
        self.fopt.graph_mode = True
        graph = torch.cuda.CUDAGraph()
        n0 = _nl.LAUNCH_COUNT["n"]
        with torch.cuda.graph(graph):
            g["loss"] = self._eager_step(criterion, s_in, s_tg)
        g["launches"] = _nl.LAUNCH_COUNT["n"] - n0
        _nl.LAUNCH_COUNT["n"] = n0               # captured, not executed
        self.fopt.step_count -= 1  # capture does not execute; the step is counted at replay
        g["graph"] = graph

    def _make_set(self, data, batch_size=None, labels=None):
        if isinstance(data, FeatureSet):
            return data
        if hasattr(data, "to_feature_set"):
            return data.to_feature_set()
        if isinstance(data, (tuple, list)) and len(data) == 2 and labels is None:
            return FeatureSet.numpy(data[0], data[1])
        if isinstance(data, (np.ndarray, torch.Tensor)):
            return FeatureSet.numpy(data, labels)
        if isinstance(data, torch.utils.data.DataLoader):
            return FeatureSet.pytorch_dataloader(data)
        raise TypeError(f"unsupported data type {type(data)}")

    def train(self, train_set, criterion, end_trigger=None, checkpoint_trigger=None, validation_set=None,
              validation_method=None, batch_size=32, shard=True, log_every=None):
        """`batch_size` is the GLOBAL batch (reference semantics); each rank steps batch_size / world."""
        ctx = self.ctx
        criterion = to_loss(criterion).to(self.device)
        self._last_criterion = criterion
        self.prepare()
        self._ensure_run_dir()
        end_trigger = end_trigger or MaxEpoch(self.state["epoch"])
        if not isinstance(end_trigger, ZooTrigger):
            raise TypeError("end_trigger must be a ZooTrigger")
        fs = self._make_set(train_set)
        if shard and ctx.world_size > 1:
            fs = fs.shard(ctx.rank, ctx.world_size)
        if batch_size % ctx.world_size:
            raise ValueError(f"batch_size {batch_size} must be divisible by the number of workers {ctx.world_size}")
        local_bs = batch_size // ctx.world_size
        metrics = [to_metric(m, criterion) for m in (validation_method or [])]
        conf = ctx.conf
        if self._dropper is None and float(conf.get("zoo.drop.percentage", 0) or 0) > 0:
            self.set_drop_module_property(float(conf.get("zoo.drop.percentage")), float(conf.get("zoo.drop.maxPercentage", conf.get("zoo.drop.percentage"))),
                                          int(conf.get("zoo.drop.computeThresholdBatchSize", 100)), int(conf.get("zoo.drop.warmupIterationNum", 200)),
                                          int(conf.get("zoo.drop.microBatches", 4)))
        retry_budget, retry_window = int(conf.get("zoo.failure.retryTimes", 5)), float(conf.get("zoo.failure.retryTimeInterval", 120))
        failures, window_start = 0, time.time()
        self.state["numSlice"] = fs.num_of_slice()
        wd_timeout = float(os.environ.get("ZOO_WATCHDOG_TIMEOUT", conf.get("zoo.watchdog.timeout", 0)) or 0)
        if wd_timeout > 0 and self.watchdog is None:      # heartbeats + hang detection across the ranks of the box
            from ..utils.watchdog import RankWatchdog

            self.watchdog = RankWatchdog(ctx.world_size, ctx.rank, timeout=wd_timeout).start()
        # A restarted job (torchrun --max-restarts after a watchdog abort, or conf zoo.failure.autoResume) continues from the
        # newest checkpoint of an earlier run directory instead of starting over.
        resume = str(conf.get("zoo.failure.autoResume", os.environ.get("ZOO_AUTO_RESUME", "auto")))
        if resume != "0" and not getattr(self, "_resumed", False):
            self._resumed = True
            if resume == "1" or int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0") or 0) > 0:
                prev = self.find_resume_checkpoint()
                if prev is not None:
                    log.warning("resuming from %s", prev)
                    self.load_checkpoint(prev)
        # Ranks agree at every step boundary on "go" or "somebody failed" (zoo.utils.agreement); only needed when a retry is
        # possible at all, i.e. when checkpoints are being written.
        from ..utils.agreement import PeerFailure, StepAgreement

        self._agree = None
        if ctx.world_size > 1 and self.ckpt_dir is not None and checkpoint_trigger is not None and str(conf.get("zoo.failure.agree", "1")) != "0":
            self._agree_gen = getattr(self, "_agree_gen", 0) + 1
            apath = StepAgreement.default_path(f"agree-{self._agree_gen}")
            if ctx.rank == 0:
                StepAgreement.reset(apath, ctx.world_size)
            ctx.barrier()
            self._agree = StepAgreement(ctx.world_size, ctx.rank, path=apath, timeout=float(conf.get("zoo.failure.agreeTimeout", 120)))
        try:
            while True:
                try:
                    self._train_loop(fs, criterion, end_trigger, checkpoint_trigger, validation_set, metrics, local_bs,
                                     batch_size, log_every)
                    break
                except (ValueError, TypeError, KeyboardInterrupt):
                    raise
                except Exception as e:  # retry from the newest checkpoint (Topology.scala:1209-1258)
                    if ctx.world_size > 1:
                        if self._agree is None:
                            raise
                        if not isinstance(e, PeerFailure):
                            self._agree.report_failure()     # peers waiting at this step boundary learn about it and recover too
                    if self.latest_checkpoint() is None:
                        raise
                    now = time.time()
                    if now - window_start > retry_budget * retry_window:
                        failures, window_start = 0, now
                    failures += 1
                    if failures > retry_budget:
                        raise
                    log.warning("training failed with %r; retry %d/%d from %s", e, failures, retry_budget, self.latest_checkpoint())
                    self.load_checkpoint()
        finally:
            if self._agree is not None:
                self._agree.close()
                self._agree = None
        if self.async_loss and getattr(self, "_final_loss_dev", None) is not None:
            self.state["Loss"] = float(self._final_loss_dev)
            self._final_loss_dev = None
        if self.train_summary is not None:
            self.train_summary.writer.flush()
        return self

    def _maybe_inject_fault(self):
        spec = os.environ.get("ZOO_FAULT_INJECT")
        if spec:
            r, it = spec.split(":")
            if int(r) == self.ctx.rank and int(it) == self.state["neval"]:
                os.environ.pop("ZOO_FAULT_INJECT")
                raise RuntimeError(f"injected fault at rank {r} iteration {it}")

    def _train_loop(self, fs, criterion, end_trigger, ckpt_trigger, val_set, metrics, local_bs, global_bs, log_every):
        st, dev = self.state, self.device
        self.model.train()
        n_local = fs.size()
        iters_per_pass = max(1, (n_local // fs.num_of_slice()) // local_bs) if fs.num_of_slice() > 1 else max(1, n_local // local_bs)
        while not end_trigger(st):
            t_epoch = time.time()
            it = Prefetcher(fs.batches(local_bs, train=True, drop_last=True), dev)
            seen, t_last, loss_dev = 0, time.time(), None
            for batch in it:
                self._maybe_inject_fault()
                if getattr(self, "_agree", None) is not None:
                    self._agree.check_in()          # every rank starts this step, or all of them recover together
                self.fopt.epoch = st["epoch"]
                loss_dev = self.train_step(criterion, batch)
                if self.watchdog is not None:
                    self.watchdog.beat(st["neval"])
                if self.on_step is not None:
                    self.on_step(st["neval"], loss_dev)
                st["neval"] += 1
                seen += batch.size
                st["recordsProcessedThisEpoch"] += batch.size * self.ctx.world_size
                every = log_every or max(1, iters_per_pass // 4)
                if (st["neval"] - 1) % every == 0 or seen + local_bs > n_local:
                    st["Loss"] = self._read_loss(loss_dev)  # the only host sync of the loop
                    dt = max(time.time() - t_last, 1e-9)
                    thr = every * global_bs / dt if (st["neval"] - 1) % every == 0 else seen * self.ctx.world_size / max(time.time() - t_epoch, 1e-9)
                    self.last_throughput = thr
                    t_last = time.time()
                    log.info("[Epoch %d %d/%d][Iteration %d] Trained %d records. Throughput is %.1f records/second. Loss is %.6f",
                             st["epoch"], st["recordsProcessedThisEpoch"], n_local * self.ctx.world_size, st["neval"] - 1,
                             global_bs, thr, st["Loss"])
                    if self.train_summary is not None:
                        s = st["neval"] - 1
                        self.train_summary.add_scalar("Loss", st["Loss"], s).add_scalar("Throughput", thr, s)
                        self.train_summary.add_scalar("LearningRate", self.fopt._lr_now[0], s)
                if ckpt_trigger is not None and not isinstance(ckpt_trigger, EveryEpoch) and ckpt_trigger(st):
                    self._validate_and_checkpoint(val_set, metrics, criterion, local_bs)
                if end_trigger(st) and not isinstance(end_trigger, MaxEpoch):
                    break
            if loss_dev is not None:
                self._final_loss_dev = loss_dev
                if not self.async_loss:
                    st["Loss"] = float(loss_dev)    # (async_loss: read once when fit returns -- no pipeline drain per pass)
            # one pass over the (slice of the) data is done
            if fs.num_of_slice() > 1:
                st["currentSlice"] = fs.current_slice
                if fs.current_slice % fs.num_of_slice() == 0:
                    st["epoch"] += 1
                    st["recordsProcessedThisEpoch"] = 0
            else:
                st["epoch"] += 1
                st["recordsProcessedThisEpoch"] = 0
            if ckpt_trigger is not None and isinstance(ckpt_trigger, EveryEpoch):
                if ckpt_trigger(st) or True:
                    self._validate_and_checkpoint(val_set, metrics, criterion, local_bs)

    def _read_loss(self, loss_dev):
        """Device -> host read of the step's loss.  With `async_loss` the copy of step i is issued
        asynchronously into pinned memory and the value of step i-1 is returned, so the host never
        waits for the step it has just launched."""
        if not (self.async_loss and loss_dev.is_cuda):
            return float(loss_dev)
        if self._loss_ring is None:
            self._loss_ring = {"buf": torch.zeros(2, dtype=torch.float32).pin_memory(), "ev": [None, None], "i": 0}
        r = self._loss_ring
        i = r["i"]
        r["buf"][i:i + 1].copy_(loss_dev.detach().reshape(1).float(), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        r["ev"][i] = ev
        j = 1 - i
        r["i"] = j
        if r["ev"][j] is not None:
            r["ev"][j].synchronize()
            return float(r["buf"][j])
        return float("nan")

    def _validate_and_checkpoint(self, val_set, metrics, criterion, local_bs):
        if val_set is not None and metrics:
            res = self.evaluate(val_set, metrics, batch_size=local_bs * self.ctx.world_size, criterion=criterion)
            if res:
                self.state["score"] = res[0].value
            if self.val_summary is not None:
                for r in res:
                    self.val_summary.add_scalar(r.name, r.value, self.state["neval"] - 1)
            self.metrics_history.append((self.state["neval"] - 1, res))
            self.model.train()
        self.save_checkpoint()

    # ------------------------------------------------------- evaluate/predict --
    @torch.no_grad()
    def evaluate(self, val_set, validation_method, batch_size=32, criterion=None, shard=True):
        ctx = self.ctx
        self.model.to(self.device)
        criterion = to_loss(criterion).to(self.device) if criterion is not None else self._last_criterion
        metrics = [to_metric(m, criterion) for m in validation_method]
        for m in metrics:
            m.reset()
        fs = self._make_set(val_set)
        if shard and ctx.world_size > 1:
            fs = fs.shard(ctx.rank, ctx.world_size)
        was_training = self.model.training
        self.model.eval()
        local_bs = max(1, batch_size // ctx.world_size)
        for batch in Prefetcher(fs.batches(local_bs, train=False), self.device):
            x = batch.inputs
            out = self.model(*x) if isinstance(x, tuple) else self.model(x)
            loss = None
            if criterion is not None and any(type(m).__name__ == "Loss" for m in metrics):
                loss = criterion(out.float() if torch.is_tensor(out) else out, batch.targets)
            for m in metrics:
                m.update(out, batch.targets, loss)
        if ctx.world_size > 1:
            import torch.distributed as dist

            for m in metrics:
                t = m.state().to(self.device if ctx.backend == "nccl" else "cpu")
                dist.all_reduce(t)
                m.load(t.cpu())
        self.model.train(was_training)
        return [m.result() for m in metrics]

    @torch.no_grad()
    def predict(self, data, batch_size=32, to_numpy=True):
        self.model.to(self.device)
        fs = self._make_set(data)
        was_training = self.model.training
        self.model.eval()
        outs = []
        for batch in Prefetcher(fs.batches(batch_size, train=False), self.device):
            x = batch.inputs
            o = self.model(*x) if isinstance(x, tuple) else self.model(x)
            outs.append(o.float().cpu() if torch.is_tensor(o) else tuple(t.float().cpu() for t in o))
        self.model.train(was_training)
        if not outs:
            return np.zeros((0,))
        if torch.is_tensor(outs[0]):
            r = torch.cat(outs)
            return r.numpy() if to_numpy else r
        r = tuple(torch.cat([o[i] for o in outs]) for i in range(len(outs[0])))
        return tuple(t.numpy() for t in r) if to_numpy else r

    def close(self):
        for s in (self.train_summary, self.val_summary):
            if s is not None:
                s.close()
        if self._dp is not None:
            self._dp.close()
        if self.watchdog is not None:
            self.watchdog.stop()
            self.watchdog = None
