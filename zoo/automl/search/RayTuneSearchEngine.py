"""Hyper-parameter search engine.

Reference: pyzoo/zoo/automl/search/RayTuneSearchEngine.py:28-458 -- `compile` builds a trainable closure
(feature transform -> model.fit_eval per iteration -> report the reward metric -> save a zip of the best
state), `run` hands it to `ray.tune.run(config=search_space, stop=..., num_samples=...,
resources_per_trial=...)`, `get_best_trials(k)` sorts by the reward metric.

Ray Tune is not part of this image.  The engine expands the search space itself (GridSearch = cartesian
product, RandomSample = one draw per sample, `num_samples` repetitions; a `BayesOpt` recipe's box ranges are
sampled uniformly and then refined around the best point) and runs the trials on the box: inline, or one per
worker of an initialised `RayContext` pool -- one GPU per worker, which is the `resources_per_trial` of the
reference.  The class name is kept so user code does not change.
"""
from __future__ import annotations

import copy
import itertools
import os
import tempfile
import uuid

import numpy as np

from ..common.metrics import Evaluator
from ..common.util import save_zip
from ..config.recipe import GridSearch, RandomSample, convert_bayes_configs
from .abstract import SearchEngine, TrialOutput


class _Spec:
    class _Cfg(dict):
        __getattr__ = dict.__getitem__

    def __init__(self, config):
        self.config = _Spec._Cfg(config)


def expand_search_space(space, num_samples, rng=None):
    """-> list of concrete configs.  Dependent RandomSamples are resolved after the independent ones."""
    grid_keys = [k for k, v in space.items() if isinstance(v, GridSearch)]
    grids = list(itertools.product(*[space[k].values for k in grid_keys])) if grid_keys else [()]
    out = []
    for _ in range(max(1, int(num_samples))):
        for combo in grids:
            cfg = {k: v for k, v in space.items() if not isinstance(v, (GridSearch, RandomSample))}
            cfg.update(dict(zip(grid_keys, combo)))
            pending = {k: v for k, v in space.items() if isinstance(v, RandomSample)}
            for _round in range(len(pending) + 1):
                for k in list(pending):
                    try:
                        val = pending[k].func(_Spec(cfg))
                    except (KeyError, AttributeError):
                        continue             # depends on a value not sampled yet
                    cfg[k] = val.item() if isinstance(val, np.generic) else val
                    del pending[k]
                if not pending:
                    break
            if pending:
                raise ValueError(f"could not resolve search-space entries {sorted(pending)}")
            out.append(cfg)
    return out


def _run_trial(payload):
    """One trial: returns (reward, config, zip path).  Top-level so worker processes can import it."""
    (config, input_df, validation_df, ft, model_create, metric, training_iteration, reward_target, mc, out_dir, fixed) = payload
    import torch

    np.random.seed(abs(hash(str(sorted(config.items(), key=str)))) % (2**31))
    torch.manual_seed(0)
    config = convert_bayes_configs({**config, **(fixed or {})})
    ft = copy.deepcopy(ft)
    model = model_create()
    x, y = ft.fit_transform(input_df, **config)
    val = None
    if validation_df is not None and (isinstance(validation_df, list) or len(validation_df) > 0):
        val = ft.transform(validation_df, is_train=True)
    mode = Evaluator.get_metric_mode(metric)
    best = None
    path = os.path.join(out_dir, f"trial_{uuid.uuid4().hex[:12]}.zip")
    for it in range(int(training_iteration)):
        reward = model.fit_eval(x, y, validation_data=val, mc=mc, metric=metric, **config)
        better = best is None or (reward > best if mode == "max" else reward < best)
        if better:
            best = reward
            save_zip(path, ft, model, config)
        if reward_target is not None:
            signed = reward if mode == "max" else -reward
            if signed >= reward_target:
                break
    return best, config, path


class RayTuneSearchEngine(SearchEngine):
    def __init__(self, logs_dir="", resources_per_trial=None, name="", remote_dir=None):
        self.logs_dir = logs_dir or tempfile.mkdtemp(prefix="zoo_automl_")
        self.resources_per_trial, self.name, self.remote_dir = resources_per_trial, name, remote_dir
        self.trials, self.results = [], []
        self.compiled = None

    def compile(self, input_df, model_create_func, search_space, recipe=None, feature_transformers=None, future_seq_len=1,
                validation_df=None, mc=False, metric="mse"):
        Evaluator.check_metric(metric)
        rt = recipe.runtime_params() if recipe is not None else {"training_iteration": 1, "num_samples": 1}
        self.metric, self.mode = metric, Evaluator.get_metric_mode(metric)
        os.makedirs(self.logs_dir, exist_ok=True)
        fixed = recipe.fixed_params() if recipe is not None else None
        algo = recipe.search_algorithm() if recipe is not None else None
        if algo == "BayesOpt":      # box ranges -> uniform samples
            space = {k: (RandomSample(lambda spec, lo=v[0], hi=v[1]: np.random.uniform(lo, hi)) if isinstance(v, tuple) else v)
                     for k, v in search_space.items()}
        else:
            space = search_space
        self.trials = expand_search_space(space, rt.get("num_samples", 1))
        self.compiled = dict(input_df=input_df, validation_df=validation_df, ft=feature_transformers, model_create=model_create_func,
                             training_iteration=rt.get("training_iteration", 1), reward_target=rt.get("reward_metric"), mc=mc,
                             fixed=fixed, bayes=algo == "BayesOpt", box=search_space if algo == "BayesOpt" else None)
        return self

    def _payload(self, cfg):
        c = self.compiled
        return (cfg, c["input_df"], c["validation_df"], c["ft"], c["model_create"], self.metric, c["training_iteration"],
                c["reward_target"], c["mc"], self.logs_dir, c["fixed"])

    def _execute(self, configs):
        from ...ray.raycontext import RayContext

        ctx = RayContext._active
        payloads = [self._payload(c) for c in configs]
        if ctx is not None and ctx.initialized and not ctx.local_mode and len(payloads) > 1:
            return ctx.map(_run_trial, payloads)
        return [_run_trial(p) for p in payloads]

    def run(self):
        if self.compiled is None:
            raise RuntimeError("compile(...) must be called before run()")
        self.results = self._execute(self.trials)
        if self.compiled["bayes"] and self.results:     # one refinement round around the incumbent
            best_cfg = self._sorted()[0][1]
            box = self.compiled["box"]
            refine = []
            for _ in range(max(1, len(self.trials) // 2)):
                c = {}
                for k, v in box.items():
                    if isinstance(v, tuple):
                        raw = best_cfg.get(k[:-6] if k.endswith("_float") else k, (v[0] + v[1]) / 2)
                        raw = raw if isinstance(raw, (int, float)) else (v[0] + v[1]) / 2
                        c[k] = float(np.clip(raw + np.random.normal(0, 0.1 * (v[1] - v[0])), v[0], v[1]))
                    else:
                        c[k] = v
                for k in box:
                    if k.startswith("bayes_feature_"):
                        c[k] = 1.0 if k[len("bayes_feature_"):] in best_cfg.get("selected_features", []) else c[k]
                refine.append(c)
            self.results += self._execute(refine)
        return self.results

    def _sorted(self):
        ok = [r for r in self.results if r[0] is not None and np.isfinite(r[0])]
        return sorted(ok, key=lambda r: r[0], reverse=self.mode == "max")

    def get_best_trials(self, k=1):
        return [TrialOutput(config=c, model_path=p) for _, c, p in self._sorted()[:k]]

    def test_run(self):
        """One trial inline (reference RayTuneSearchEngine.test_run with a mock reporter)."""
        return _run_trial(self._payload(self.trials[0]))
