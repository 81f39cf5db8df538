"""Recipes: search space + runtime parameters of a hyper-parameter search
(reference: pyzoo/zoo/automl/config/recipe.py:22-518).

The reference expresses spaces with `ray.tune.sample_from` / `grid_search`; Ray is not part of this image,
so the two primitives are defined here: `RandomSample(fn)` (fn receives a `spec` whose `.config` holds the
values sampled so far, so parameters can depend on others) and `GridSearch(values)`.  Recipe names, their
constructor arguments and the hyper-parameter keys are the reference's.
"""
from __future__ import annotations

import json
from abc import ABCMeta, abstractmethod

import numpy as np


class RandomSample:
    def __init__(self, func):
        self.func = func


class GridSearch:
    def __init__(self, values):
        self.values = list(values)


def sample_from(func):
    return RandomSample(func)


def grid_search(values):
    return GridSearch(values)


class Recipe(metaclass=ABCMeta):
    def __init__(self):
        self.training_iteration = 1
        self.num_samples = 1
        self.reward_metric = None

    @abstractmethod
    def search_space(self, all_available_features):
        pass

    def runtime_params(self):
        cfg = {"training_iteration": self.training_iteration, "num_samples": self.num_samples}
        if self.reward_metric is not None:
            cfg["reward_metric"] = self.reward_metric
        return cfg

    def fixed_params(self):
        return None

    def search_algorithm_params(self):
        return None

    def search_algorithm(self):
        return None

    def scheduler_params(self):
        return None


class SmokeRecipe(Recipe):
    """One random sample, one epoch: a smoke test."""

    def search_space(self, all_available_features):
        return {"selected_features": all_available_features, "model": "LSTM",
                "lstm_1_units": RandomSample(lambda spec: np.random.choice([32, 64], size=1)[0]),
                "dropout_1": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)),
                "lstm_2_units": RandomSample(lambda spec: np.random.choice([32, 64], size=1)[0]),
                "dropout_2": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)),
                "lr": 0.001, "batch_size": 1024, "epochs": 1, "past_seq_len": 2}


class MTNetSmokeRecipe(Recipe):
    def search_space(self, all_available_features):
        return {"selected_features": all_available_features, "model": "MTNet", "lr": 0.001, "batch_size": 16, "epochs": 1,
                "dropout": 0.2, "time_step": RandomSample(lambda spec: np.random.choice([3, 4], size=1)[0]), "filter_size": 2,
                "long_num": RandomSample(lambda spec: np.random.choice([3, 4], size=1)[0]),
                "ar_size": RandomSample(lambda spec: np.random.choice([2, 3], size=1)[0]),
                "past_seq_len": RandomSample(lambda spec: (spec.config.long_num + 1) * spec.config.time_step)}


class TCNSmokeRecipe(Recipe):
    """North-star extension: the TCN forecaster (kernel K7)."""

    def search_space(self, all_available_features):
        return {"selected_features": all_available_features, "model": "TCN", "lr": 0.001, "batch_size": 64, "epochs": 1,
                "dropout": 0.1, "kernel_size": 3, "hidden_units": RandomSample(lambda spec: np.random.choice([16, 30], size=1)[0]),
                "levels": 3, "past_seq_len": RandomSample(lambda spec: np.random.choice([8, 12], size=1)[0])}


class PastSeqParamHandler:
    """`look_back` may be one value or a (min, max) range (recipe.py:109-153)."""

    @staticmethod
    def get_past_seq_config(look_back):
        if isinstance(look_back, (tuple, list)) and len(look_back) == 2 and all(isinstance(v, int) for v in look_back):
            if look_back[1] < 2:
                raise ValueError(f"The max look back value should be at least 2. But found {look_back[1]}")
            lo = max(look_back[0], 2)
            return RandomSample(lambda spec: np.random.randint(lo, look_back[1] + 1))
        if isinstance(look_back, int):
            if look_back < 2:
                raise ValueError(f"look back value should not be smaller than 2. Current value is {look_back}")
            return look_back
        raise ValueError(f"look back is {look_back}.\n look_back should be either a tuple with 2 int values: (min_len, max_len) "
                         "or a single int")


def _feature_subset(all_available_features):
    return RandomSample(lambda spec: json.dumps(list(np.random.choice(
        all_available_features, size=np.random.randint(low=min(3, len(all_available_features)), high=len(all_available_features) + 1),
        replace=False))))


class GridRandomRecipe(Recipe):
    """Grid over units/batch sizes, random over the rest (recipe.py:156-213)."""

    def __init__(self, num_rand_samples=1, look_back=2, epochs=5, training_iteration=10):
        super().__init__()
        self.num_samples, self.training_iteration = num_rand_samples, training_iteration
        self.past_seq_config = PastSeqParamHandler.get_past_seq_config(look_back)
        self.epochs = epochs

    def search_space(self, all_available_features):
        return {"selected_features": _feature_subset(all_available_features), "model": RandomSample(lambda spec: np.random.choice(["LSTM", "Seq2seq"])),
                "lstm_1_units": GridSearch([16, 32]), "dropout_1": 0.2, "lstm_2_units": GridSearch([16, 32]),
                "dropout_2": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)),
                "latent_dim": GridSearch([32, 64]), "dropout": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)),
                "lr": 0.001, "batch_size": GridSearch([32, 64]), "epochs": self.epochs, "past_seq_len": self.past_seq_config}


class LSTMGridRandomRecipe(Recipe):
    def __init__(self, num_rand_samples=1, epochs=5, training_iteration=10, look_back=2, lstm_1_units=(16, 32, 64, 128),
                 lstm_2_units=(16, 32, 64), batch_size=(32, 64)):
        super().__init__()
        self.num_samples, self.training_iteration, self.epochs = num_rand_samples, training_iteration, epochs
        self.past_seq_config = PastSeqParamHandler.get_past_seq_config(look_back)
        self.lstm_1_units_config, self.lstm_2_units_config = GridSearch(lstm_1_units), GridSearch(lstm_2_units)
        self.batch_size = list(batch_size)

    def search_space(self, all_available_features):
        return {"selected_features": _feature_subset(all_available_features), "model": "LSTM",
                "lstm_1_units": self.lstm_1_units_config, "dropout_1": 0.2, "lstm_2_units": self.lstm_2_units_config,
                "dropout_2": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)), "lr": 0.001,
                "batch_size": GridSearch(self.batch_size), "epochs": self.epochs, "past_seq_len": self.past_seq_config}


class MTNetGridRandomRecipe(Recipe):
    def __init__(self, num_rand_samples=1, epochs=5, training_iteration=10, time_step=(3, 4), long_num=(3, 4), cnn_height=(2, 3),
                 cnn_hid_size=(32, 50, 100), ar_size=(2, 3), batch_size=(32, 64)):
        super().__init__()
        self.num_samples, self.training_iteration, self.epochs = num_rand_samples, training_iteration, epochs
        self.time_step, self.long_num, self.cnn_height = list(time_step), list(long_num), list(cnn_height)
        self.cnn_hid_size, self.ar_size, self.batch_size = list(cnn_hid_size), list(ar_size), list(batch_size)

    def search_space(self, all_available_features):
        return {"selected_features": _feature_subset(all_available_features), "model": "MTNet", "lr": 0.001,
                "batch_size": GridSearch(self.batch_size), "epochs": self.epochs,
                "cnn_dropout": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)),
                "rnn_dropout": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)),
                "time_step": GridSearch(self.time_step), "long_num": GridSearch(self.long_num), "ar_size": GridSearch(self.ar_size),
                "cnn_height": GridSearch(self.cnn_height), "cnn_hid_size": GridSearch(self.cnn_hid_size),
                "past_seq_len": RandomSample(lambda spec: (spec.config.long_num + 1) * spec.config.time_step)}


class RandomRecipe(Recipe):
    def __init__(self, num_rand_samples=1, look_back=2, epochs=5, reward_metric=-0.05, training_iteration=10):
        super().__init__()
        self.num_samples, self.training_iteration, self.epochs = num_rand_samples, training_iteration, epochs
        self.reward_metric = reward_metric
        self.past_seq_config = PastSeqParamHandler.get_past_seq_config(look_back)

    def search_space(self, all_available_features):
        return {"selected_features": _feature_subset(all_available_features), "model": RandomSample(lambda spec: np.random.choice(["LSTM", "Seq2seq"])),
                "lstm_1_units": RandomSample(lambda spec: np.random.choice([8, 16, 32, 64, 128])),
                "dropout_1": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)),
                "lstm_2_units": RandomSample(lambda spec: np.random.choice([8, 16, 32, 64, 128])),
                "dropout_2": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)),
                "latent_dim": RandomSample(lambda spec: np.random.choice([32, 64, 128, 256])),
                "dropout": RandomSample(lambda spec: np.random.uniform(0.2, 0.5)), "lr": 0.001,
                "batch_size": RandomSample(lambda spec: np.random.choice([32, 64, 1024])), "epochs": self.epochs,
                "past_seq_len": self.past_seq_config}


class BayesRecipe(Recipe):
    """Continuous ranges for a Bayesian-optimisation style search (recipe.py:423-518).  Without the `bayes_opt`
    package the engine samples the same ranges uniformly and refines around the best trial."""

    def __init__(self, num_samples=1, look_back=2, epochs=5, reward_metric=-0.05, training_iteration=5):
        super().__init__()
        self.num_samples, self.training_iteration, self.epochs, self.reward_metric = num_samples, training_iteration, epochs, reward_metric
        if isinstance(look_back, (tuple, list)) and len(look_back) == 2:
            if look_back[1] < 2:
                raise ValueError(f"The max look back value should be at least 2. But found {look_back[1]}")
            self.bayes_past_seq_config = {"past_seq_len_float": (max(look_back[0], 2), look_back[1])}
            self.fixed_past_seq_config = {}
        elif isinstance(look_back, int):
            if look_back < 2:
                raise ValueError(f"look back value should not be smaller than 2. Current value is {look_back}")
            self.bayes_past_seq_config, self.fixed_past_seq_config = {}, {"past_seq_len": look_back}
        else:
            raise ValueError(f"look back is {look_back}.\n look_back should be either a tuple with 2 int values: (min_len, max_len) or a single int")

    def search_space(self, all_available_features):
        feature_space = {f"bayes_feature_{f}": (0.3, 1) for f in all_available_features}
        other = {"lstm_1_units_float": (8, 128), "dropout_1": (0.2, 0.5), "lstm_2_units_float": (8, 128), "dropout_2": (0.2, 0.5),
                 "batch_size_float": (32, 1024)}
        return {**feature_space, **other, **self.bayes_past_seq_config}

    def fixed_params(self):
        return {"epochs": self.epochs, "model": "LSTM", "lr": 0.001, **self.fixed_past_seq_config}

    def search_algorithm_params(self):
        return {"utility_kwargs": {"kind": "ucb", "kappa": 2.5, "xi": 0.0}}

    def search_algorithm(self):
        return "BayesOpt"


def convert_bayes_configs(config):
    """`*_float` -> rounded ints, `bayes_feature_*` > 0.5 -> selected_features (common/util.py)."""
    out, sel = {}, []
    for k, v in config.items():
        if k.startswith("bayes_feature_"):
            if v >= 0.5:
                sel.append(k[len("bayes_feature_"):])
        elif k.endswith("_float"):
            out[k[:-len("_float")]] = int(round(v))
        else:
            out[k] = v
    if sel or any(k.startswith("bayes_feature_") for k in config):
        out["selected_features"] = sel
    return out
