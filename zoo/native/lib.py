"""ctypes loader for the in-tree native libraries.

``libzoo_cuda.so`` carries every hand-written sm_100a kernel behind a C ABI (declared in
``include/zoo_b200.h``); ``libzoo_rt.so`` carries the C++ runtime (stream broker, arena
allocator, record store).  On a machine with a GPU a missing or unloadable CUDA library is a
hard error -- the framework never falls back to eager PyTorch silently for an op that has a
kernel.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
_cuda = None
_rt = None

c_void_p, c_int, c_ll, c_float = C.c_void_p, C.c_int, C.c_longlong, C.c_float

# name -> argtypes (all launchers return int; 0 == ok)
_CUDA_SIGS = {
    "zoo_dense_fwd_bf16": [c_void_p, c_ll, c_void_p, c_ll, c_void_p, c_void_p, c_ll, c_int, c_int, c_int, c_int, c_void_p],
    "zoo_dense_dgrad_bf16": [c_void_p, c_ll, c_void_p, c_ll, c_void_p, c_ll, c_int, c_void_p, c_ll, c_int, c_int, c_int, c_void_p],
    "zoo_dense_fwd_bf16_bits": [c_void_p, c_ll, c_void_p, c_ll, c_void_p, c_void_p, c_ll, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p],
    "zoo_dense_dgrad_bf16_bits": [c_void_p, c_ll, c_void_p, c_ll, c_void_p, c_int, c_int, c_void_p, c_ll, c_int, c_int, c_int, c_void_p],
    "zoo_attention_fwd_bf16": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int, c_void_p],
    "zoo_attention_fwd_lse_bf16": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int, c_void_p],
    "zoo_attention_bwd_bf16": [c_void_p] * 11 + [c_int, c_int, c_int, c_int, c_float, c_int, c_void_p],
    "zoo_dense_wgrad_bias_bf16": [c_void_p, c_ll, c_void_p, c_ll, c_void_p, c_ll, c_void_p, c_int, c_int, c_int, c_void_p],
    "zoo_dense_wgrad_bf16": [c_void_p, c_ll, c_void_p, c_ll, c_void_p, c_ll, c_int, c_int, c_int, c_void_p],
    "zoo_embedding_gather": [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_ll, c_void_p, c_ll, c_int,
                             c_int, c_int, c_void_p, c_ll, c_void_p, c_void_p, c_void_p],
    "zoo_embedding_scratch_init": [c_void_p, c_ll, c_void_p],
    "zoo_embedding_sparse_update": [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_ll, c_void_p, c_ll, c_float, c_ll,
                                    c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p],
    "zoo_embedding_sparse_update_w": [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_ll, c_void_p, c_ll, c_float, c_ll,
                                      c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p],
    "zoo_embedding_dense_grad": [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_ll, c_void_p, c_ll, c_float, c_ll,
                                 c_int, c_void_p, c_void_p],
    "zoo_dense_opt_step": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_void_p, c_int, c_void_p],
    "zoo_hyper_step": [c_void_p, c_void_p, c_void_p],
    "zoo_sumsq": [c_void_p, c_ll, c_void_p, c_void_p, c_int, c_void_p],
    "zoo_clip_l2": [c_void_p, c_ll, c_void_p, c_float, c_void_p],
    "zoo_clip_const": [c_void_p, c_ll, c_float, c_float, c_void_p],
    "zoo_colsum_bf16": [c_void_p, c_ll, c_int, c_int, c_void_p, c_void_p],
    "zoo_cast_f32_to_bf16": [c_void_p, c_void_p, c_ll, c_void_p],
    "zoo_cast_bf16_to_f32": [c_void_p, c_void_p, c_ll, c_void_p],
    "zoo_softmax_nll": [c_void_p, c_ll, c_void_p, c_int, c_int, c_int, c_int, c_ll, c_int, c_void_p, c_float, c_float,
                        c_float, c_void_p, c_void_p, c_void_p, c_ll, c_void_p, c_ll, c_void_p],
    "zoo_sigmoid_bce": [c_void_p, c_void_p, c_ll, c_int, c_int, c_float, c_float, c_float, c_void_p, c_void_p, c_void_p],
}


class NativeError(RuntimeError):
    pass


def cuda_lib_path() -> Path:
    return _HERE / "libzoo_cuda.so"


def rt_lib_path() -> Path:
    return _HERE / "libzoo_rt.so"


def _bind(lib, sigs):
    for name, argtypes in sigs.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue  # optional symbol (older build); callers fail loudly when they need it
        fn.argtypes = argtypes
        fn.restype = c_int


def cuda_lib():
    """Return the loaded CUDA kernel library (building it if the .so is absent and nvcc exists)."""
    global _cuda
    if _cuda is not None:
        return _cuda
    path = cuda_lib_path()
    if not path.exists():
        if os.environ.get("ZOO_NO_AUTOBUILD") != "1" and Path("/usr/local/cuda/bin/nvcc").exists():
            from .build import build_cuda

            build_cuda()
        if not path.exists():
            raise NativeError(f"{path} is missing: run `python -m zoo.native.build` (needs nvcc)")
    lib = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
    _bind(lib, _CUDA_SIGS)
    lib.zoo_last_error.restype = C.c_char_p
    lib.zoo_sumsq_scratch_bytes.restype = c_ll
    _cuda = lib
    return lib


def register_cuda_sigs(sigs: dict):
    """Let op modules add signatures for symbols they own (keeps this file small)."""
    _CUDA_SIGS.update(sigs)
    if _cuda is not None:
        _bind(_cuda, sigs)


LAUNCH_COUNT = {"n": 0}
_KERNELS_PER_CALL = {"zoo_embedding_sparse_update": 3, "zoo_embedding_sparse_update_w": 3, "zoo_sumsq": 1}


def check(rc: int, what: str = ""):
    """Raise on a non-zero status; also counts kernel launches (each launcher = 1 kernel unless listed)."""
    LAUNCH_COUNT["n"] += _KERNELS_PER_CALL.get(what, 1)
    if rc != 0:
        msg = cuda_lib().zoo_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed (code {rc}): {msg}")


def have_cuda_lib() -> bool:
    return cuda_lib_path().exists()


def cuda_available() -> bool:
    """True when torch sees a GPU; the native kernels are then mandatory."""
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def rt_lib():
    global _rt
    if _rt is not None:
        return _rt
    path = rt_lib_path()
    if not path.exists():
        if os.environ.get("ZOO_NO_AUTOBUILD") != "1":
            from .build import build_runtime

            build_runtime()
        if not path.exists():
            raise NativeError(f"{path} is missing: run `python -m zoo.native.build`")
    _rt = C.CDLL(str(path))
    return _rt


def ptr(t):
    """Device/host pointer of a tensor (or None)."""
    return None if t is None else t.data_ptr()


def stream_ptr():
    import torch

    return torch.cuda.current_stream().cuda_stream
