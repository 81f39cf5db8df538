"""Remote file systems (hdfs://, s3://, s3a://, gs://, ...) behind one small interface.

Reference: zoo/.../common/Utils.scala:35-278 and utils/File.scala read and write models, checkpoints and data through Hadoop's
`FileSystem` API (HDFS, S3 through s3a); pyzoo/zoo/xshard/pandas/preprocessing.py:26-58 lists and opens files on HDFS / S3 with
pyarrow / boto3.  Here every non-local URI is resolved with `pyarrow.fs.FileSystem.from_uri` (libhdfs for HDFS -- it needs a Hadoop
client installation on the machine --, the AWS SDK built into pyarrow for S3, GCS likewise); schemes can also be registered by
hand (`register_filesystem`), which is how the tests mount a directory as `mock://` and how a site can plug in its own store.
Local paths and `file://` never touch pyarrow.
"""
from __future__ import annotations

import os
import shutil

_REGISTRY = {}      # scheme -> (pyarrow.fs.FileSystem, path prefix inside it)


def register_filesystem(scheme, fs, root=""):
    """Serve `scheme://...` from `fs` (any pyarrow.fs.FileSystem); paths are taken relative to `root`."""
    _REGISTRY[scheme.rstrip(":/")] = (fs, root.rstrip("/"))


def unregister_filesystem(scheme):
    _REGISTRY.pop(scheme.rstrip(":/"), None)


def is_local(path):
    path = str(path)
    return "://" not in path or path.startswith("file://")


def local_path(path):
    path = str(path)
    return path[len("file://"):] if path.startswith("file://") else path


def resolve(uri):
    """uri -> (pyarrow filesystem, path inside it).  Raises a clear error when the scheme's client is not usable here."""
    uri = str(uri)
    scheme, rest = uri.split("://", 1)
    if scheme in _REGISTRY:
        fs, root = _REGISTRY[scheme]
        return fs, (root + "/" + rest.lstrip("/")) if root else rest
    import pyarrow.fs as pafs

    if scheme == "s3a" or scheme == "s3n":      # Hadoop's names for S3
        uri = "s3://" + rest
    try:
        return pafs.FileSystem.from_uri(uri)
    except Exception as e:  # libhdfs missing, no credentials / endpoint, unknown scheme, ...
        raise OSError(f"cannot open the {scheme}:// file system for {uri!r}: {e}.  HDFS needs a Hadoop client (libhdfs, CLASSPATH) on "
                      "this machine, S3 / GCS need credentials and network access; other stores can be plugged in with "
                      "zoo.util.filesystem.register_filesystem(scheme, pyarrow_filesystem)") from e


def exists(path):
    if is_local(path):
        return os.path.exists(local_path(path))
    import pyarrow.fs as pafs

    fs, p = resolve(path)
    return fs.get_file_info(p).type != pafs.FileType.NotFound


def isdir(path):
    if is_local(path):
        return os.path.isdir(local_path(path))
    import pyarrow.fs as pafs

    fs, p = resolve(path)
    return fs.get_file_info(p).type == pafs.FileType.Directory


def ls(path, recursive=False, suffix=None):
    """Files under `path` (a file or a directory), sorted, as full URIs of the same scheme."""
    if is_local(path):
        p = local_path(path)
        if os.path.isfile(p):
            out = [p]
        else:
            out = []
            for root, _, files in os.walk(p):
                out += [os.path.join(root, f) for f in files]
                if not recursive:
                    break
        out = sorted(out)
    else:
        import pyarrow.fs as pafs

        scheme = str(path).split("://", 1)[0]
        fs, p = resolve(path)
        info = fs.get_file_info(p)
        if info.type == pafs.FileType.NotFound:
            raise FileNotFoundError(path)
        if info.type == pafs.FileType.File:
            infos = [info]
        else:
            infos = [i for i in fs.get_file_info(pafs.FileSelector(p, recursive=recursive)) if i.type == pafs.FileType.File]
        root = _REGISTRY.get(scheme, (None, ""))[1]
        strip = (lambda q: q[len(root):].lstrip("/")) if root else (lambda q: q)
        out = sorted(f"{scheme}://{strip(i.path)}" for i in infos)
    return [f for f in out if suffix is None or f.endswith(suffix)]


def open_file(path, mode="rb"):
    """A file object for `path`; remote files are pyarrow streams (binary)."""
    if is_local(path):
        return open(local_path(path), mode)
    fs, p = resolve(path)
    if "r" in mode:
        return fs.open_input_file(p)
    if "a" in mode:
        return fs.open_append_stream(p)
    parent = p.rsplit("/", 1)[0] if "/" in p else ""
    if parent:
        fs.create_dir(parent, recursive=True)
    return fs.open_output_stream(p)


def makedirs(path):
    if is_local(path):
        os.makedirs(local_path(path), exist_ok=True)
    else:
        fs, p = resolve(path)
        fs.create_dir(p, recursive=True)


def get(remote, local, over_write=False):
    """Copy `remote` (any scheme) to the local file `local`."""
    local = local_path(local)
    if os.path.exists(local) and not over_write:
        raise FileExistsError(local)
    if is_local(remote):
        shutil.copy(local_path(remote), local)
        return local
    with open_file(remote, "rb") as src, open(local, "wb") as dst:
        shutil.copyfileobj(src, dst, 16 << 20)
    return local


def put(local, remote, over_write=False):
    """Copy the local file `local` to `remote` (any scheme)."""
    if exists(remote) and not over_write:
        raise FileExistsError(remote)
    if is_local(remote):
        os.makedirs(os.path.dirname(local_path(remote)) or ".", exist_ok=True)
        shutil.copy(local_path(local), local_path(remote))
        return remote
    with open(local_path(local), "rb") as src, open_file(remote, "wb") as dst:
        shutil.copyfileobj(src, dst, 16 << 20)
    return remote


def remove(path):
    if is_local(path):
        p = local_path(path)
        shutil.rmtree(p) if os.path.isdir(p) else os.remove(p)
    else:
        import pyarrow.fs as pafs

        fs, p = resolve(path)
        if fs.get_file_info(p).type == pafs.FileType.Directory:
            fs.delete_dir(p)
        else:
            fs.delete_file(p)
