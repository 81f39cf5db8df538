"""In-tree native build: every CUDA kernel and C++ runtime component of zoo-b200.

`python -m zoo.native.build` (or `__graft_entry__.build()`) compiles

* ``csrc/**/*.cu``        -> ``zoo/native/libzoo_cuda.so``  (sm_100a only, C ABI, loaded with ctypes)
* ``csrc/runtime/*.cpp``  -> ``zoo/native/libzoo_rt.so``    (stream broker, arena allocator, record
  store, TFRecord/CRC writer -- the C++ runtime that replaces the reference's JNI pieces,
  SURVEY.md section 2.3 rows N4/N9)

The shared objects are built in-tree so they travel with a repository snapshot; nothing is
JIT-compiled into a per-user cache.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
OUT = Path(__file__).resolve().parent
OBJ = ROOT / "build" / "obj"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--use_fast_math",
    "-I", str(ROOT / "include"), "-I", str(CSRC),
]
CXX = os.environ.get("CXX", "g++")
CXX_FLAGS = ["-O2", "-msse4.2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-pthread", "-I", str(ROOT / "include"), "-I", str(CSRC)]


def _digest(paths, flags) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(str(p).encode())
        h.update(Path(p).read_bytes())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def _run(cmd, verbose):
    if verbose:
        print(" ".join(map(str, cmd)), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("native build failed: " + " ".join(map(str, cmd[:3])) + " ...")
    if verbose and (r.stdout or r.stderr):
        print(r.stdout + r.stderr)


def _headers():
    return list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")) + list((ROOT / "include").glob("*.h"))


def build_cuda(force=False, verbose=False, ptxas_info=False) -> Path:
    srcs = sorted(CSRC.rglob("*.cu"))
    target = OUT / "libzoo_cuda.so"
    stamp = OUT / ".libzoo_cuda.sha256"
    flags = NVCC_FLAGS + (["-Xptxas", "-v"] if ptxas_info else [])
    dig = _digest(srcs + _headers(), flags)
    if not force and target.exists() and stamp.exists() and stamp.read_text() == dig:
        return target
    OBJ.mkdir(parents=True, exist_ok=True)
    objs = []

    def compile_one(src):
        obj = OBJ / (src.relative_to(CSRC).as_posix().replace("/", "_") + ".o")
        _run([NVCC, *flags, "-c", str(src), "-o", str(obj)], verbose or ptxas_info)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs) or 1)) as ex:
        objs = list(ex.map(compile_one, srcs))
    _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(target), *map(str, objs)], verbose)
    stamp.write_text(dig)
    return target


def build_runtime(force=False, verbose=False) -> Path:
    srcs = sorted((CSRC / "runtime").glob("*.cpp"))
    target = OUT / "libzoo_rt.so"
    stamp = OUT / ".libzoo_rt.sha256"
    if not srcs:
        return target
    dig = _digest(srcs + _headers(), CXX_FLAGS)
    if not force and target.exists() and stamp.exists() and stamp.read_text() == dig:
        return target
    _run([CXX, *CXX_FLAGS, "-shared", "-o", str(target), *map(str, srcs), "-ldl", "-lrt"], verbose)
    stamp.write_text(dig)
    return target


def build_all(force=False, verbose=False):
    return build_cuda(force, verbose), build_runtime(force, verbose)


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("-v", "--verbose", action="store_true")
    ap.add_argument("--ptxas-info", action="store_true", help="print registers/spills/smem per kernel")
    a = ap.parse_args()
    if a.ptxas_info:
        build_cuda(True, True, True)
    else:
        print(*build_all(a.force, a.verbose), sep="\n")
