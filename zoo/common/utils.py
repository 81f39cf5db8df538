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
    """Files under a local, hdfs:// or s3:// path (reference: Utils.listPaths, common/Utils.scala:52-73)."""
    from ..util import filesystem as zfs

    return zfs.ls(path, recursive)


def get_remote_file_to_local(remote_path, local_path, over_write=False):
    """Utils.getRemoteFileToLocal (common/Utils.scala:249-264): any scheme zoo.util.filesystem resolves."""
    from ..util import filesystem as zfs

    return zfs.get(remote_path, local_path, over_write)


def put_local_file_to_remote(local_path, remote_path, over_write=False):
    from ..util import filesystem as zfs

    return zfs.put(local_path, remote_path, over_write)


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
    """Run `save_func(local_path)`; for a remote `path` the file is written to a temporary local file and uploaded (the reference
    does the same dance through Hadoop's FileSystem, utils/File.scala)."""
    from ..util import filesystem as zfs

    if is_local_path(path):
        return save_func(zfs.local_path(path), **kwargs)
    import tempfile

    with tempfile.TemporaryDirectory() as d:
        tmp = os.path.join(d, os.path.basename(path) or "file")
        save_func(tmp, **kwargs)
        zfs.put(tmp, path, over_write=True)


def load_from_file(load_func, path):
    from ..util import filesystem as zfs

    if is_local_path(path):
        return load_func(zfs.local_path(path))
    import tempfile

    with tempfile.TemporaryDirectory() as d:
        tmp = os.path.join(d, os.path.basename(path) or "file")
        zfs.get(path, tmp, over_write=True)
        return load_func(tmp)
