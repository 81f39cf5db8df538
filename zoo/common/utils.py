"""Common utilities (reference: pyzoo/zoo/common/utils.py:25-166): `JTensor`, `Sample`, remote file get/put,
`set_core_number`, and `callZooFunc` -- in the reference a py4j call into a JVM `PythonZoo*` object; here a
name -> Python callable registry so extension code written against `callZooFunc(bigdl_type, name, *args)`
keeps working."""
from __future__ import annotations

import os
import shutil

import numpy as np
import torch

from ..feature.common import Sample  # noqa: F401  (re-export: zoo.common.utils.Sample in the reference)


class JTensor:
    """ndarray + shape (+ sparse indices) container (the py4j wire format of the reference)."""

    def __init__(self, storage, shape, bigdl_type="float", indices=None):
        self.storage = np.asarray(storage, dtype=np.float32 if bigdl_type == "float" else np.float64).reshape(-1)
        self.shape = np.asarray(shape, dtype=np.int32)
        self.indices = None if indices is None else np.asarray(indices, dtype=np.int32)
        self.bigdl_type = bigdl_type

    @classmethod
    def from_ndarray(cls, a_ndarray, bigdl_type="float"):
        if a_ndarray is None:
            return None
        a = np.asarray(a_ndarray)
        return cls(a, a.shape if a.shape else (a.size,), bigdl_type)

    @classmethod
    def sparse(cls, a_ndarray, i_ndarray, shape, bigdl_type="float"):
        """values + [ndim, nnz] indices + dense shape."""
        return cls(a_ndarray, shape, bigdl_type, np.asarray(i_ndarray).reshape(-1))

    def to_ndarray(self):
        if self.indices is None:
            return self.storage.reshape(tuple(self.shape))
        nd = len(self.shape)
        idx = self.indices.reshape(nd, -1)
        out = np.zeros(tuple(self.shape), dtype=self.storage.dtype)
        out[tuple(idx)] = self.storage
        return out

    def to_tensor(self):
        if self.indices is None:
            return torch.from_numpy(self.to_ndarray())
        nd = len(self.shape)
        return torch.sparse_coo_tensor(torch.from_numpy(self.indices.reshape(nd, -1).astype(np.int64)), torch.from_numpy(self.storage),
                                       tuple(int(s) for s in self.shape))

    def __repr__(self):
        return f"JTensor: storage: {self.storage}, shape: {self.shape}" + ("" if self.indices is None else f", indices {self.indices}")


_ZOO_FUNCS = {}


def register_zoo_func(name, fn=None):
    if fn is None:
        return lambda f: register_zoo_func(name, f)
    _ZOO_FUNCS[name] = fn
    return fn


def callZooFunc(bigdl_type, name, *args):
    """Dispatch by name (reference: resolves the method on the registered JVM `PythonZoo*` creators)."""
    if name not in _ZOO_FUNCS:
        raise Exception(f"Cannot find function: {name}")
    return _ZOO_FUNCS[name](*args)


def _is_remote(path):
    return "://" in path and not path.startswith("file://")


def get_file_list(path, recursive=False):
    path = path[len("file://"):] if path.startswith("file://") else path
    if _is_remote(path):
        raise NotImplementedError("remote file systems (hdfs://, s3://) need clients that are not on this box")
    if os.path.isfile(path):
        return [path]
    out = []
    for root, _, files in os.walk(path):
        out += [os.path.join(root, f) for f in sorted(files)]
        if not recursive:
            break
    return out


def get_remote_file_to_local(remote_path, local_path, over_write=False):
    src = remote_path[len("file://"):] if remote_path.startswith("file://") else remote_path
    if _is_remote(src):
        raise NotImplementedError("remote file systems (hdfs://, s3://) need clients that are not on this box")
    if os.path.exists(local_path) and not over_write:
        raise FileExistsError(local_path)
    shutil.copy(src, local_path)
    return local_path


def put_local_file_to_remote(local_path, remote_path, over_write=False):
    return get_remote_file_to_local(local_path, remote_path, over_write)


def set_core_number(num):
    torch.set_num_threads(max(1, int(num)))


register_zoo_func("setCoreNumber", set_core_number)
register_zoo_func("getRemoteFileToLocal", get_remote_file_to_local)
register_zoo_func("putLocalFileToRemote", put_local_file_to_remote)
register_zoo_func("listPaths", get_file_list)


def is_local_path(path):
    return "://" not in path or path.startswith("file://")


def append_suffix(prefix, path):
    """`prefix` + the file extension of `path`."""
    return prefix + os.path.splitext(path)[1]


def convert_to_safe_path(input_path, follow_symlinks=True):
    return os.path.realpath(input_path) if follow_symlinks else os.path.abspath(input_path)


def to_list_of_numpy(elements):
    if isinstance(elements, np.ndarray):
        return [elements]
    if np.isscalar(elements):
        return [np.array(elements)]
    if not isinstance(elements, list):
        raise ValueError(f"Wrong type: {type(elements)}")
    return [e if isinstance(e, np.ndarray) else np.array(e) for e in elements]


def save_file(save_func, path, **kwargs):
    if not is_local_path(path):
        raise NotImplementedError("remote file systems need clients that are not on this box")
    save_func(path[len("file://"):] if path.startswith("file://") else path, **kwargs)


def load_from_file(load_func, path):
    if not is_local_path(path):
        raise NotImplementedError("remote file systems need clients that are not on this box")
    return load_func(path[len("file://"):] if path.startswith("file://") else path)
