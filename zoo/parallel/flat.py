"""Flat parameter arena + fused optimizer (the B200 replacement of the reference's
``PytorchModel`` JNI flat-buffer view and BigDL ``AllReduceParameter``).

The reference copies the whole model three times per iteration between libtorch and a JVM
float[] (``updateWeightNative`` / ``getGradientNative`` + ``arraycopy``,
zoo/.../pipeline/api/net/TorchNet.scala:86-139) and then runs the optimizer slice-wise after a
BlockManager reduce (Topology.scala:1128-1206).  Here the parameters *are* the flat buffer:

* every trainable parameter is re-pointed into one contiguous fp32 arena, its gradient into a
  second arena (autograd accumulates in place, fused kernels add with `red.global`);
* optimizer state (m, v) and an optional bf16 shadow of the weights are parallel arenas;
* one kernel launch per parameter group applies the optimizer, writes the bf16 shadow and
  zeroes the gradient (``csrc/ops/optim.cu``), or -- across GPUs -- the fused
  reduce-scatter + optimizer + all-gather kernel of ``csrc/comm/rs_opt_ag.cu``.

Per-sub-module optimizers (``Estimator(model, optim_methods={"widePart": Ftrl, "deepPart":
Adagrad})``, Estimator.scala:65-68) become parameter groups = contiguous arena ranges.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch
import torch.nn as nn

from ..native import lib as _nl

OPT_IDS = {"sgd": 0, "adam": 1, "adagrad": 2, "adam_zoo": 3, "adamw_bert": 4, "ftrl": 5, "rmsprop": 6, "adamw": 7}
HYPER_LEN = 16
H_LR, H_B1, H_B2, H_EPS, H_WD, H_STEP, H_GSCALE, H_X1, H_X2, H_X3 = range(10)
_ALIGN = 8  # elements; keeps every tensor 32-byte aligned in fp32 and 16-byte aligned in bf16


@dataclass
class OptConfig:
    """Optimizer of one parameter group, normalised from torch.optim.* or zoo optimizers."""
    name: str = "adam"
    lr: float = 1e-3
    b1: float = 0.9          # adam beta1 / sgd momentum
    b2: float = 0.999        # adam beta2 / rmsprop alpha
    eps: float = 1e-8
    weight_decay: float = 0.0
    x1: float = 0.0          # sgd dampening / ftrl l1
    x2: float = 0.0          # sgd nesterov / ftrl l2
    x3: float = -0.5         # ftrl lr_power
    init_accum: float = 0.0  # adagrad / ftrl initial accumulator
    schedule: object = None  # callable(step:int, epoch:int) -> lr multiplier or absolute lr via .lr_at

    @staticmethod
    def from_torch(opt: torch.optim.Optimizer, group_index=0) -> "OptConfig":
        g = opt.param_groups[group_index]
        n = type(opt).__name__.lower()
        if n == "adam":
            return OptConfig("adam", g["lr"], g["betas"][0], g["betas"][1], g["eps"], g.get("weight_decay", 0.0))
        if n == "adamw":
            return OptConfig("adamw", g["lr"], g["betas"][0], g["betas"][1], g["eps"], g.get("weight_decay", 0.0))
        if n == "sgd":
            return OptConfig("sgd", g["lr"], g.get("momentum", 0.0), 0.0, 0.0, g.get("weight_decay", 0.0),
                             x1=g.get("dampening", 0.0), x2=1.0 if g.get("nesterov", False) else 0.0)
        if n == "adagrad":
            return OptConfig("adagrad", g["lr"], 0.0, 0.0, g.get("eps", 1e-10), g.get("weight_decay", 0.0),
                             init_accum=g.get("initial_accumulator_value", 0.0))
        if n == "rmsprop":
            return OptConfig("rmsprop", g["lr"], 0.0, g.get("alpha", 0.99), g.get("eps", 1e-8), g.get("weight_decay", 0.0))
        raise ValueError(f"no fused kernel for optimizer {type(opt).__name__}; use Adam/AdamW/SGD/Adagrad/RMSprop")

    def hyper_vector(self, lr=None, step=0.0, gscale=1.0):
        h = [0.0] * HYPER_LEN
        h[H_LR], h[H_B1], h[H_B2], h[H_EPS], h[H_WD] = (self.lr if lr is None else lr), self.b1, self.b2, self.eps, self.weight_decay
        h[H_STEP], h[H_GSCALE], h[H_X1], h[H_X2], h[H_X3] = step, gscale, self.x1, self.x2, self.x3
        return h

    @property
    def needs_m(self):
        return self.name != "sgd" or self.b1 != 0.0

    @property
    def needs_v(self):
        return self.name in ("adam", "adam_zoo", "adamw_bert", "ftrl", "adamw")


@dataclass
class ParamGroup:
    name: str
    start: int
    end: int
    cfg: OptConfig
    params: list = field(default_factory=list)


def _round_up(n, a):
    return (n + a - 1) // a * a


class FlatParamArena:
    """Re-points parameters (and their .grad) into contiguous fp32 arenas."""

    def __init__(self, named_params, device=None, with_bf16_shadow=None, allocator=None, pad_to=1):
        named_params = [(n, p) for n, p in named_params if p.requires_grad]
        if not named_params:
            raise ValueError("no trainable parameters")
        self.device = device or named_params[0][1].device
        self.names = [n for n, _ in named_params]
        self.params = [p for _, p in named_params]
        self.align = max(pad_to, _ALIGN)   # every tensor (hence every group boundary) starts on this many elements
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += _round_up(p.numel(), self.align)
        self.numel = _round_up(off, self.align)
        alloc = allocator or (lambda n, dtype: torch.zeros(n, dtype=dtype, device=self.device))
        self.flat_p = alloc(self.numel, torch.float32)
        self.flat_g = alloc(self.numel, torch.float32)
        if with_bf16_shadow is None:
            with_bf16_shadow = self.device.type == "cuda"
        self.flat_p16 = alloc(self.numel, torch.bfloat16) if with_bf16_shadow else None
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                n = p.numel()
                # A channels_last conv weight keeps its physical NHWC order inside the arena (the optimizer is
                # elementwise, so order is irrelevant to it, while cuDNN gets the layout it was promised).
                cl = p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last) and not p.is_contiguous()
                src = p.detach().permute(0, 2, 3, 1) if cl else p.detach()
                self.flat_p[o:o + n].copy_(src.reshape(-1).to(self.device, torch.float32))
                view = (lambda f: f[o:o + n].view(p.shape[0], p.shape[2], p.shape[3], p.shape[1]).permute(0, 3, 1, 2)) if cl \
                    else (lambda f: f[o:o + n].view(p.shape))
                p.data = view(self.flat_p)
                g = view(self.flat_g)
                p.grad = g
                p._zoo_grad = g
            if self.flat_p16 is not None:
                self.flat_p16.copy_(self.flat_p)

    def range_of(self, param):
        i = next(k for k, q in enumerate(self.params) if q is param)
        return self.offsets[i], self.offsets[i] + param.numel()

    def shadow_of(self, param):
        if self.flat_p16 is None:
            return None
        a, b = self.range_of(param)
        return self.flat_p16[a:b].view(param.shape)

    def reattach_grads(self):
        """Restore .grad views if user code dropped them (e.g. zero_grad(set_to_none=True))."""
        for p, o in zip(self.params, self.offsets):
            if p.grad is None or p.grad.data_ptr() != self.flat_g.data_ptr() + 4 * o:
                p.grad = p._zoo_grad


def _torch_opt_update(cfg: OptConfig, p, g, m, v, lr, step, gscale):
    """CPU restatement of csrc/ops/optim_math.cuh (same formulas, fp32)."""
    g = g * gscale
    n = cfg.name
    if n in ("adam", "adamw", "adam_zoo"):
        if n == "adam":
            g = g + cfg.weight_decay * p
        if n == "adamw":
            p.mul_(1.0 - lr * cfg.weight_decay)
        m.mul_(cfg.b1).add_(g, alpha=1 - cfg.b1)
        v.mul_(cfg.b2).addcmul_(g, g, value=1 - cfg.b2)
        t = max(step, 1.0)
        bc1, bc2 = 1 - cfg.b1 ** t, 1 - cfg.b2 ** t
        if n == "adam_zoo":
            p.addcdiv_(m, v.sqrt() + cfg.eps, value=-lr * math.sqrt(bc2) / bc1)
        else:
            p.addcdiv_(m / bc1, (v / bc2).sqrt() + cfg.eps, value=-lr)
    elif n == "adamw_bert":
        m.mul_(cfg.b1).add_(g, alpha=1 - cfg.b1)
        v.mul_(cfg.b2).addcmul_(g, g, value=1 - cfg.b2)
        p.sub_(lr * (m / (v.sqrt() + cfg.eps) + cfg.weight_decay * p))
    elif n == "adagrad":
        g = g + cfg.weight_decay * p
        m.addcmul_(g, g)
        p.addcdiv_(g, m.sqrt() + cfg.eps, value=-lr)
    elif n == "rmsprop":
        g = g + cfg.weight_decay * p
        m.mul_(cfg.b2).addcmul_(g, g, value=1 - cfg.b2)
        p.addcdiv_(g, m.sqrt() + cfg.eps, value=-lr)
    elif n == "ftrl":
        l1, l2, lrp = cfg.x1, cfg.x2, cfg.x3
        acc_new = m + g * g
        sigma = (acc_new.pow(-lrp) - m.pow(-lrp)) / lr
        v.add_(g - sigma * p)
        quad = acc_new.pow(-lrp) / lr + 2 * l2
        p.copy_(torch.where(v.abs() > l1, (torch.sign(v) * l1 - v) / quad, torch.zeros_like(p)))
        m.copy_(acc_new)
    elif n == "sgd":
        g = g + cfg.weight_decay * p
        if cfg.b1 != 0.0:
            if step <= 1.0:
                m.copy_(g)
            else:
                m.mul_(cfg.b1).add_(g, alpha=1 - cfg.x1)
            g = g + cfg.b1 * m if cfg.x2 != 0.0 else m
        p.sub_(lr * g)
    else:
        raise ValueError(n)


class FusedOptimizer:
    """Optimizer over a :class:`FlatParamArena`; one fused pass per parameter group."""

    def __init__(self, arena: FlatParamArena, groups: list[ParamGroup]):
        self.arena = arena
        self.groups = groups
        dev = arena.device
        self.m = torch.zeros_like(arena.flat_p) if any(g.cfg.needs_m for g in groups) else None
        self.v = torch.zeros_like(arena.flat_p) if any(g.cfg.needs_v for g in groups) else None
        for g in groups:
            if g.cfg.init_accum and self.m is not None:
                self.m[g.start:g.end].fill_(g.cfg.init_accum)
        self.step_count = 0
        self.epoch = 0
        self.hyper_host = torch.tensor([g.cfg.hyper_vector() for g in groups], dtype=torch.float32)
        if dev.type == "cuda":
            self.hyper_host = self.hyper_host.pin_memory()
        self.hyper_dev = self.hyper_host.to(dev)
        self.hyper_stage = None
        self.grad_scale = 1.0
        self._lr_now = [g.cfg.lr for g in groups]
        self._dirty = True
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self._norm_scratch = None
        self.clip_l2 = None
        self.clip_const = None
        self.graph_mode = False   # inside a captured step: hyper-parameters are re-read from pinned memory
        self._side = None         # side stream of `begin_step`
        self._begun = None        # event recorded there once this step's hyper-parameters are live on the device

    def pre_graph_replay(self):
        """Host-side part of a step whose device part is replayed from a CUDA graph."""
        self.step_count += 1
        self._apply_schedules()
        for i in range(len(self.groups)):
            self.hyper_host[i, H_LR] = self._lr_now[i]
            self.hyper_host[i, H_GSCALE] = self.grad_scale

    # -- configuration ---------------------------------------------------------
    @staticmethod
    def build(model: nn.Module, default_cfg: OptConfig, per_module: dict | None = None, device=None, allocator=None,
              pad_to=1):
        """Create arena + groups.  `per_module` maps a sub-module name to its OptConfig
        (reference: Estimator(model, optim_methods: Map[subModuleName -> OptimMethod]))."""
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        per_module = per_module or {}

        def cfg_key(pname):
            best = None
            for k in per_module:
                if pname == k or pname.startswith(k + "."):
                    if best is None or len(k) > len(best):
                        best = k
            return best

        order = sorted(range(len(named)), key=lambda i: (str(cfg_key(named[i][0])), i))
        named = [named[i] for i in order]
        arena = FlatParamArena(named, device=device, allocator=allocator, pad_to=pad_to)
        groups, cur = [], None
        for (n, p), off in zip(named, arena.offsets):
            k = cfg_key(n)
            if cur is None or cur.name != str(k):
                cur = ParamGroup(str(k), off, off, per_module[k] if k is not None else default_cfg)
                groups.append(cur)
            cur.end = off + _round_up(p.numel(), arena.align)
            cur.params.append(p)
        groups[-1].end = arena.numel
        return FusedOptimizer(arena, groups)

    def set_lr(self, lr, group=None):
        for i, g in enumerate(self.groups):
            if group is None or group == i or group == g.name:
                self._lr_now[i] = float(lr)
        self._dirty = True

    def set_grad_scale(self, s):
        self.grad_scale = float(s)
        self._dirty = True

    def set_gradient_clipping_by_l2_norm(self, clip_norm):
        self.clip_l2, self.clip_const = float(clip_norm), None

    def set_constant_gradient_clipping(self, lo, hi):
        self.clip_const, self.clip_l2 = (float(lo), float(hi)), None

    def clear_gradient_clipping(self):
        self.clip_l2 = self.clip_const = None

    def _sync_hyper(self):
        """Push host-side hyper-parameters (lr after schedules, gradient scale) to the device: one
        contiguous pinned->device copy into a staging buffer; `zoo_hyper_step` then merges it into
        the live array while advancing the device-side step.  Returns True when a refresh is due."""
        if not self._dirty and not self.graph_mode:
            return False
        for i, g in enumerate(self.groups):
            self.hyper_host[i, H_LR] = self._lr_now[i]
            self.hyper_host[i, H_GSCALE] = self.grad_scale
        if self.arena.device.type == "cuda":
            if self.hyper_stage is None:
                self.hyper_stage = torch.zeros_like(self.hyper_dev)
            self.hyper_stage.copy_(self.hyper_host, non_blocking=True)
        self._dirty = False
        return True

    def _apply_schedules(self):
        for i, g in enumerate(self.groups):
            sch = g.cfg.schedule
            if sch is not None:
                lr = sch(self.step_count, self.epoch, g.cfg.lr)
                if lr != self._lr_now[i]:
                    self._lr_now[i] = float(lr)
                    self._dirty = True

    # -- the step ---------------------------------------------------------------
    def clip_gradients(self):
        """Clipping acts on the AVERAGED gradient (the reference divides the aggregated gradient by the replica count before
        its clipping processors run).  `flat_g` holds the SUM over ranks and 1/world is applied later inside the optimizer
        kernel (`grad_scale`), so the thresholds are expressed on the sum: bounds and norm limit divided by `grad_scale`."""
        a = self.arena
        gs = self.grad_scale if self.grad_scale > 0 else 1.0
        if self.clip_const is not None:
            lo, hi = self.clip_const[0] / gs, self.clip_const[1] / gs
            if a.device.type == "cuda":
                _nl.check(_nl.cuda_lib().zoo_clip_const(a.flat_g.data_ptr(), a.numel, lo, hi, _nl.stream_ptr()), "zoo_clip_const")
            else:
                a.flat_g.clamp_(lo, hi)
        elif self.clip_l2 is not None:
            if a.device.type == "cuda":
                lib = _nl.cuda_lib()
                if self._norm_scratch is None:
                    self._norm_scratch = torch.zeros(int(lib.zoo_sumsq_scratch_bytes()), dtype=torch.uint8, device=a.device)
                _nl.check(lib.zoo_sumsq(a.flat_g.data_ptr(), a.numel, self._norm_scratch.data_ptr(), self.sumsq.data_ptr(), 0,
                                        _nl.stream_ptr()), "zoo_sumsq")
                _nl.check(lib.zoo_clip_l2(a.flat_g.data_ptr(), a.numel, self.sumsq.data_ptr(), self.clip_l2 / gs, _nl.stream_ptr()),
                          "zoo_clip_l2")
            else:
                norm = a.flat_g.norm()
                if norm > self.clip_l2 / gs:
                    a.flat_g.mul_((self.clip_l2 / gs) / (norm + 1e-6))

    def begin_step(self):
        """Optional, at the START of a training step (before forward): the pinned->device copy of the hyper-parameters and the
        `zoo_hyper_step` launches do not depend on the step's gradients, so they run on a side stream next to forward / backward
        -- in a captured step a parallel branch of the graph -- instead of as ~10 us of serial tail (copy node, kernel, two
        launch gaps) before the optimizer kernel.  `step()` joins the branch."""
        a = self.arena
        if a.device.type != "cuda" or self._begun is not None:
            return
        self.step_count += 1
        self._apply_schedules()
        if self._side is None:
            self._side = torch.cuda.Stream(a.device)
        cur = torch.cuda.current_stream(a.device)
        self._side.wait_stream(cur)
        with torch.cuda.stream(self._side):
            refresh = self._sync_hyper()
            lib, st = _nl.cuda_lib(), _nl.stream_ptr()
            for i in range(len(self.groups)):
                src = self.hyper_stage[i].data_ptr() if refresh else None
                _nl.check(lib.zoo_hyper_step(self.hyper_dev[i].data_ptr(), src, st), "zoo_hyper_step")
            ev = torch.cuda.Event()
            ev.record(self._side)
        self._begun = ev

    def step(self, zero_grad=True):
        a = self.arena
        begun, self._begun = self._begun, None
        if begun is not None:
            torch.cuda.current_stream(a.device).wait_event(begun)
            refresh = False
        else:
            self.step_count += 1
            self._apply_schedules()
            refresh = self._sync_hyper()
        self.clip_gradients()
        if a.device.type == "cuda":
            lib, st = _nl.cuda_lib(), _nl.stream_ptr()
            for i, g in enumerate(self.groups):
                n = g.end - g.start
                hp = self.hyper_dev[i].data_ptr()
                src = self.hyper_stage[i].data_ptr() if refresh else None
                if begun is None:
                    _nl.check(lib.zoo_hyper_step(hp, src, st), "zoo_hyper_step")
                _nl.check(lib.zoo_dense_opt_step(
                    a.flat_p[g.start:].data_ptr(), a.flat_g[g.start:].data_ptr(),
                    self.m[g.start:].data_ptr() if (self.m is not None and g.cfg.needs_m) else None,
                    self.v[g.start:].data_ptr() if (self.v is not None and g.cfg.needs_v) else None,
                    a.flat_p16[g.start:].data_ptr() if a.flat_p16 is not None else None,
                    n, OPT_IDS[g.cfg.name], hp, int(zero_grad), st), "zoo_dense_opt_step")
        else:
            with torch.no_grad():
                for i, g in enumerate(self.groups):
                    sl = slice(g.start, g.end)
                    m = self.m[sl] if self.m is not None else None
                    v = self.v[sl] if self.v is not None else None
                    _torch_opt_update(g.cfg, a.flat_p[sl], a.flat_g[sl], m, v, self._lr_now[i], float(self.step_count),
                                      self.grad_scale)
                if zero_grad:
                    a.flat_g.zero_()

    def zero_grad(self):
        self.arena.flat_g.zero_()

    # -- checkpointing ----------------------------------------------------------
    def state_dict(self):
        return {"step": self.step_count, "epoch": self.epoch, "lr": list(self._lr_now),
                "m": None if self.m is None else self.m.detach().cpu(),
                "v": None if self.v is None else self.v.detach().cpu(),
                "groups": [(g.name, g.start, g.end, g.cfg.name) for g in self.groups]}

    def load_state_dict(self, sd):
        self.step_count, self.epoch = int(sd["step"]), int(sd.get("epoch", 0))
        self._lr_now = list(sd["lr"])
        if sd["m"] is not None and self.m is not None:
            self.m.copy_(sd["m"])
        if sd["v"] is not None and self.v is not None:
            self.v.copy_(sd["v"])
        self.hyper_host[:, H_STEP] = float(self.step_count)
        self.hyper_dev[:, H_STEP] = float(self.step_count)
        self._dirty = True
