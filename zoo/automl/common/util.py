"""File helpers of the AutoML stack (reference: pyzoo/zoo/automl/common/util.py): config JSON round trip with
numpy types, zip / unzip of a pipeline directory, train/val/test split."""
from __future__ import annotations

import json
import os
import shutil
import tempfile
import zipfile

import numpy as np


class NumpyEncoder(json.JSONEncoder):
    def default(self, o):
        if isinstance(o, np.integer):
            return int(o)
        if isinstance(o, np.floating):
            return float(o)
        if isinstance(o, np.ndarray):
            return o.tolist()
        if isinstance(o, (np.bool_,)):
            return bool(o)
        return super().default(o)


def save_config(file_path, config, replace=False):
    os.makedirs(os.path.dirname(os.path.abspath(file_path)), exist_ok=True)
    old = {}
    if os.path.isfile(file_path) and not replace:
        with open(file_path) as f:
            old = json.load(f)
    old.update(config)
    with open(file_path, "w") as f:
        json.dump(old, f, cls=NumpyEncoder)


def load_config(file_path):
    with open(file_path) as f:
        return json.load(f)


def save_zip(file, feature_transformers=None, model=None, config=None):
    """Zip {model file, config.json (incl. scaler state)} into `file` (pipeline/time_sequence.py:188-221)."""
    d = tempfile.mkdtemp()
    try:
        model_path, config_path = os.path.join(d, "weights_tune.h5"), os.path.join(d, "all.json")
        if feature_transformers is not None:
            feature_transformers.save(config_path, replace=True)
        if model is not None:
            model.save(model_path, config_path)
        if config is not None:
            save_config(config_path, config)
        os.makedirs(os.path.dirname(os.path.abspath(file)), exist_ok=True)
        with zipfile.ZipFile(file, "w") as z:
            for f in os.listdir(d):
                z.write(os.path.join(d, f), f)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    return file


def restore_zip(file, feature_transformers=None, model=None, config=None):
    d = tempfile.mkdtemp()
    try:
        with zipfile.ZipFile(file) as z:
            z.extractall(d)
        model_path, config_path = os.path.join(d, "weights_tune.h5"), os.path.join(d, "all.json")
        all_config = load_config(config_path)
        if config:
            all_config.update(config)
        if feature_transformers is not None:
            feature_transformers.restore(**all_config)
        if model is not None:
            model.restore(model_path, **all_config)
        return all_config
    finally:
        shutil.rmtree(d, ignore_errors=True)


def split_input_df(input_df, val_split_ratio=0, test_split_ratio=0.1, ts_col="datetime", overlap=0):
    """Chronological train/val/test split (util.py `split_input_df`)."""
    n = len(input_df)
    df = input_df.sort_values(ts_col).reset_index(drop=True) if ts_col in input_df else input_df
    n_test, n_val = int(n * test_split_ratio), int(n * val_split_ratio)
    n_train = n - n_test - n_val
    train = df.iloc[:n_train]
    val = df.iloc[max(n_train - overlap, 0):n_train + n_val]
    test = df.iloc[max(n_train + n_val - overlap, 0):]
    return train.reset_index(drop=True), val.reset_index(drop=True), test.reset_index(drop=True)


def convert_bayes_configs(config):
    from ..config.recipe import convert_bayes_configs as _c

    return _c(config)


# reference names (util.py): save/restore are the zip helpers above; HDFS variants need a client that is not here
save, restore = save_zip, restore_zip


def process(command, fail_fast=False, timeout=120):
    from ...ray.process import session_execute

    return session_execute(command, fail_fast=fail_fast, timeout=timeout)


def _join(base, *parts):
    return "/".join([str(base).rstrip("/")] + [str(q).strip("/") for q in parts])


def get_remote_list(dir_in):
    """Names of the files under a remote directory (hdfs:// s3:// gs:// file://), through `zoo.util.filesystem`
    (reference: util.py:193-207 shells out to `hadoop fs -ls`)."""
    from ...util import filesystem as zfs

    return [os.path.basename(q.rstrip("/")) for q in zfs.ls(dir_in)]


def upload_ppl_hdfs(upload_dir, ckpt_name):
    """Copy the local checkpoint `ckpt_name` (a file, or every file of a directory) into `upload_dir` (reference: util.py:210-220)."""
    from ...util import filesystem as zfs

    zfs.makedirs(upload_dir)
    if os.path.isdir(ckpt_name):
        base = os.path.basename(ckpt_name.rstrip("/"))
        for root, _, files in os.walk(ckpt_name):
            rel = os.path.relpath(root, ckpt_name)
            for f in files:
                zfs.put(os.path.join(root, f), _join(upload_dir, base, *([] if rel == "." else [rel]), f), over_write=True)
    else:
        zfs.put(ckpt_name, _join(upload_dir, os.path.basename(ckpt_name)), over_write=True)


def restore_hdfs(model_path, remote_dir, feature_transformers=None, model=None, config=None):
    """Fetch the zipped pipeline `basename(model_path)` from `remote_dir` and restore it (reference: util.py:223-239)."""
    from ...util import filesystem as zfs

    name = os.path.basename(model_path)
    d = tempfile.mkdtemp(prefix="automl_save_")
    try:
        local = zfs.get(_join(remote_dir, name), os.path.join(d, name))
        return restore_zip(local, feature_transformers, model, config)
    finally:
        shutil.rmtree(d, ignore_errors=True)
