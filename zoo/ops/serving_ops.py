"""Cluster Serving batch pre/post-processing: faces of kernels K13 / K14 (``csrc/ops/serving.cu``).

`preprocess_images` packs decoded uint8 HWC images (any sizes) into one pinned buffer, copies it
once, and one kernel resizes (OpenCV INTER_LINEAR geometry), centre-crops, normalises and lays the
batch out as [B,C,H,W] -- the reference does this per image with OpenCV and a Scala copy loop
(zoo/.../serving/PreProcessing.scala:24-43, ClusterServing.scala:209-216).  `topk` is the
`PostProcessing.topN` of the reference (PostProcessing.scala:83-92) for a whole batch.
On CPU both functions compute the same thing with cv2 / torch (they are the oracles of the GPU tests).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ..native import lib as _nl

_V, _I, _L = C.c_void_p, C.c_int, C.c_longlong
_nl.register_cuda_sigs({
    "zoo_serving_preproc": [_V, _V, _V, _V, _I, _I, _I, _I, _I, _I, _V, _V, _I, _I, _I, _V],
    "zoo_serving_topk": [_V, _L, _I, _I, _I, _I, _I, _V, _V, _V],
    "zoo_jpeg_batch_create": [_I, _I, _V, _V],
    "zoo_jpeg_batch_destroy": [_V],
    "zoo_jpeg_batch_info": [_V, _V, _V, _I, _V, _V, _V],
    "zoo_jpeg_batch_decode": [_V, _V, _V, _I, _V, _V, _V, _I, _V],
})


def preprocess_reference(images, out_hw, resize_hw=None, mean=None, std=None, swap_rb=False, chw=True):
    """CPU oracle: float bilinear resize with the same geometry, crop, normalise."""
    OH, OW = out_hw
    RH, RW = resize_hw or out_hw
    outs = []
    for im in images:
        im = np.asarray(im)
        if im.ndim == 2:
            im = im[:, :, None]
        h, w, c = im.shape
        ys = (np.arange(RH) + 0.5) * (h / RH) - 0.5
        xs = (np.arange(RW) + 0.5) * (w / RW) - 0.5
        y0 = np.floor(ys).astype(int); x0 = np.floor(xs).astype(int)
        wy = ys - y0; wx = xs - x0
        wy[y0 < 0] = 0; y0[y0 < 0] = 0
        wx[x0 < 0] = 0; x0[x0 < 0] = 0
        y1 = np.minimum(y0 + 1, h - 1); x1 = np.minimum(x0 + 1, w - 1)
        wy[y0 >= h - 1] = 0; wx[x0 >= w - 1] = 0
        y0 = np.minimum(y0, h - 1); x0 = np.minimum(x0, w - 1)
        f = im.astype(np.float32)
        top = f[y0][:, x0] + wx[None, :, None] * (f[y0][:, x1] - f[y0][:, x0])
        bot = f[y1][:, x0] + wx[None, :, None] * (f[y1][:, x1] - f[y1][:, x0])
        r = top + wy[:, None, None] * (bot - top)
        cy, cx = (RH - OH) // 2, (RW - OW) // 2
        r = r[cy:cy + OH, cx:cx + OW]
        if swap_rb and c >= 3:
            r = r[:, :, [2, 1, 0] + list(range(3, c))]
        if mean is not None:
            r = r - np.asarray(mean, np.float32)[:c]
        if std is not None:
            r = r / np.asarray(std, np.float32)[:c]
        outs.append(r.transpose(2, 0, 1) if chw else r)
    return torch.from_numpy(np.stack(outs).astype(np.float32))


class ImageBatcher:
    """Reusable pinned staging + device buffers for `preprocess_images` (one H2D copy per batch)."""

    def __init__(self, device, max_bytes=64 << 20, max_batch=512):
        self.device = torch.device(device)
        self.host = torch.empty(max_bytes, dtype=torch.uint8).pin_memory()
        self.meta_host = torch.empty(max_batch * 4, dtype=torch.int64).pin_memory()
        self.dev = torch.empty(max_bytes, dtype=torch.uint8, device=self.device)
        self.meta_dev = torch.empty(max_batch * 4, dtype=torch.int64, device=self.device)
        self.max_batch = max_batch
        self._meta_ev = None        # the pinned buffers may be rewritten only after their last asynchronous copies have run

    def stage(self, images):
        B = len(images)
        if B > self.max_batch:
            raise ValueError(f"batch {B} > max_batch {self.max_batch}")
        if self._meta_ev is not None:        # small batches in quick succession: the previous H2D copies may still be queued
            self._meta_ev.synchronize()
        off, hb = 0, self.host.numpy()
        offs = self.meta_host[:B].numpy()
        hw = self.meta_host[B:2 * B].numpy().view(np.int32)    # B int64 slots viewed as 2B int32 = [B,2]
        C_ = None
        for i, im in enumerate(images):
            im = np.ascontiguousarray(im, dtype=np.uint8)
            if im.ndim == 2:
                im = im[:, :, None]
            C_ = im.shape[2] if C_ is None else C_
            if im.shape[2] != C_:
                raise ValueError("all images of a batch must have the same channel count")
            n = im.size
            if off + n > hb.size:
                raise ValueError("image batch exceeds the staging buffer")
            hb[off:off + n] = im.reshape(-1)
            offs[i] = off
            hw[2 * i], hw[2 * i + 1] = im.shape[0], im.shape[1]
            off += (n + 15) // 16 * 16
        self.dev[:off].copy_(self.host[:off], non_blocking=True)
        self.meta_dev[:2 * B].copy_(self.meta_host[:2 * B], non_blocking=True)
        if self.device.type == "cuda":
            self._meta_ev = torch.cuda.Event()
            self._meta_ev.record(torch.cuda.current_stream(self.device))
        return B, C_, off


_BATCHERS = {}


def preprocess_images(images, out_hw, resize_hw=None, mean=None, std=None, swap_rb=False, chw=True, dtype=torch.float32,
                      device=None, out=None, batcher=None):
    """images: list of uint8 HWC arrays -> tensor [B,C,OH,OW] (or [B,OH,OW,C]) on `device`."""
    device = torch.device(device) if device is not None else (torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu"))
    if device.type != "cuda":
        return preprocess_reference(images, out_hw, resize_hw, mean, std, swap_rb, chw).to(dtype)
    OH, OW = out_hw
    RH, RW = resize_hw or out_hw
    if batcher is None:
        batcher = _BATCHERS.get(device)
        if batcher is None:
            batcher = _BATCHERS[device] = ImageBatcher(device)
    B, Cc, _ = batcher.stage(images)
    shape = (B, Cc, OH, OW) if chw else (B, OH, OW, Cc)
    if out is None:
        out = torch.empty(shape, dtype=dtype, device=device)
    fa = lambda v: None if v is None else (C.c_float * 4)(*(list(np.asarray(v, np.float32).reshape(-1)) + [1.0] * 4)[:4])
    m, s = fa(mean), fa(std)
    _nl.check(_nl.cuda_lib().zoo_serving_preproc(batcher.dev.data_ptr(), batcher.meta_dev.data_ptr(),
                                                 batcher.meta_dev[B:].data_ptr(), out.data_ptr(), B, Cc, OH, OW, RH, RW,
                                                 C.cast(m, _V) if m is not None else None, C.cast(s, _V) if s is not None else None,
                                                 int(swap_rb), int(chw), int(dtype == torch.bfloat16), _nl.stream_ptr()),
              "zoo_serving_preproc")
    return out


class JpegBatchDecoder:
    """nvJPEG batched decode into device memory (csrc/ops/jpeg_batch.cu): a list of JPEG byte strings -> one packed uint8 device
    buffer of interleaved BGR images + their offsets / sizes, the layout K13 consumes.  `JpegBatchDecoder.available()` tells whether
    libnvjpeg could be loaded; the serving engine falls back to cv2 on host threads when it cannot."""

    def __init__(self, device, max_bytes=96 << 20, max_batch=512, cpu_threads=8, backend=-1):
        self.device = torch.device(device)
        lib = _nl.cuda_lib()
        h, used = C.c_void_p(), C.c_int(-1)
        with torch.cuda.device(self.device):
            _nl.check(lib.zoo_jpeg_batch_create(backend, cpu_threads, C.byref(h), C.byref(used)), "zoo_jpeg_batch_create")
        _nl.LAUNCH_COUNT["n"] -= 1          # (not a kernel launch)
        self.handle, self.backend = h, used.value
        self._by_size = {}          # padded batch size -> decoder state (nvJPEG's batched state is sized for ONE batch size)
        self._create_args = (backend, cpu_threads)
        self.dev = torch.empty(max_bytes, dtype=torch.uint8, device=self.device)
        self.meta_host = torch.empty(max_batch * 4, dtype=torch.int64).pin_memory()
        self.meta_dev = torch.empty(max_batch * 4, dtype=torch.int64, device=self.device)
        self.max_batch = max_batch
        self._meta_ev = None        # the pinned meta buffer may be rewritten only after its last asynchronous copy has run

    @staticmethod
    def available():
        try:
            return bool(_nl.cuda_lib().zoo_jpeg_available()) and torch.cuda.is_available()
        except Exception:
            return False

    def decode(self, jpegs):
        """jpegs: list of bytes.  Returns (B, offsets/hw already on the device in self.meta_dev, total bytes); the pixels are in
        self.dev.  Asynchronous on the current stream."""
        n_real = len(jpegs[1]) if isinstance(jpegs, tuple) else len(jpegs)
        if n_real > self.max_batch:
            raise ValueError(f"batch {n_real} > max_batch {self.max_batch}")
        lib = _nl.cuda_lib()
        # The engine batches continuously (whatever has arrived), so batch sizes vary; the batch is padded to the next power of two
        # with repeats of its first image (decoded into spare space) and every padded size keeps its own decoder state.
        B = 1
        while B < n_real:
            B *= 2
        B = min(B, self.max_batch)
        handle = self._by_size.get(B)
        if handle is None:
            if not self._by_size:
                handle = self.handle
            else:
                handle, used = C.c_void_p(), C.c_int(-1)
                with torch.cuda.device(self.device):
                    rc = lib.zoo_jpeg_batch_create(self.backend, self._create_args[1], C.byref(handle), C.byref(used))
                if rc != 0:
                    raise _nl.NativeError("zoo_jpeg_batch_create: " + lib.zoo_last_error().decode(errors="replace"))
            self._by_size[B] = handle
        if isinstance(jpegs, tuple):        # (buffer, offsets, lengths) from zoo.native.rt.b64_decode_batch: no per-record objects
            buf, boffs, blens = jpegs
            base = buf.ctypes.data
            addr = [base + int(o) for o in boffs] + [base + int(boffs[0])] * (B - n_real)
            size = [int(v) for v in blens] + [int(blens[0])] * (B - n_real)
            if min(size) <= 0:
                raise ValueError("a record of the batch is not valid base64")
            bufs = [buf]
        else:
            bufs = [np.frombuffer(j, dtype=np.uint8) for j in jpegs]
            addr = [b.ctypes.data for b in bufs] + [bufs[0].ctypes.data] * (B - n_real)
            size = [b.size for b in bufs] + [bufs[0].size] * (B - n_real)
        ptrs = (C.c_void_p * B)(*addr)
        lens = (C.c_size_t * B)(*size)
        ws, hs = (C.c_int * B)(), (C.c_int * B)()
        rc = lib.zoo_jpeg_batch_info(handle, ptrs, lens, B, ws, hs, None)
        if rc != 0:
            raise _nl.NativeError("zoo_jpeg_batch_info: " + lib.zoo_last_error().decode(errors="replace"))
        if self._meta_ev is not None:        # small batches in quick succession: the previous H2D copy of this buffer may still be queued
            self._meta_ev.synchronize()
        offs = self.meta_host[:B].numpy()
        hw = self.meta_host[B:2 * B].numpy().view(np.int32)
        off = 0
        for i in range(B):
            offs[i] = off
            hw[2 * i], hw[2 * i + 1] = hs[i], ws[i]
            off += (hs[i] * ws[i] * 3 + 15) // 16 * 16
        if off > self.dev.numel():
            raise ValueError("decoded batch exceeds the device staging buffer")
        coffs = (C.c_longlong * B)(*[int(o) for o in offs])
        with torch.cuda.device(self.device):
            rc = lib.zoo_jpeg_batch_decode(handle, ptrs, lens, B, self.dev.data_ptr(), coffs, ws, 1, _nl.stream_ptr())
        if rc != 0:
            raise _nl.NativeError("zoo_jpeg_batch_decode: " + lib.zoo_last_error().decode(errors="replace"))
        # offsets of the real images, then their (h, w) pairs, contiguous for K13
        if B != n_real:
            self.meta_host[n_real:n_real + n_real] = self.meta_host[B:B + n_real].clone()
        self.meta_dev[:2 * n_real].copy_(self.meta_host[:2 * n_real], non_blocking=True)
        self._meta_ev = torch.cuda.Event()
        self._meta_ev.record(torch.cuda.current_stream(self.device))
        self._keep = (getattr(self, "_keep", (None, None))[1], bufs)      # the bitstreams outlive the asynchronous decode (two generations)
        return n_real, off

    def close(self):
        handles = set(h.value for h in self._by_size.values() if h) | ({self.handle.value} if self.handle else set())
        for h in handles:
            _nl.cuda_lib().zoo_jpeg_batch_destroy(C.c_void_p(h))
        self.handle, self._by_size = None, {}


def preprocess_jpegs(jpegs, out_hw, decoder: JpegBatchDecoder, resize_hw=None, mean=None, std=None, swap_rb=False, chw=True,
                     dtype=torch.float32, out=None):
    """JPEG byte strings -> [B,3,OH,OW] on the decoder's device: nvJPEG batched decode + K13, no pixels on the host."""
    OH, OW = out_hw
    RH, RW = resize_hw or out_hw
    B, _ = decoder.decode(jpegs)
    shape = (B, 3, OH, OW) if chw else (B, OH, OW, 3)
    if out is None:
        out = torch.empty(shape, dtype=dtype, device=decoder.device)
    fa = lambda v: None if v is None else (C.c_float * 4)(*(list(np.asarray(v, np.float32).reshape(-1)) + [1.0] * 4)[:4])
    m, s = fa(mean), fa(std)
    _nl.check(_nl.cuda_lib().zoo_serving_preproc(decoder.dev.data_ptr(), decoder.meta_dev.data_ptr(), decoder.meta_dev[B:].data_ptr(),
                                                 out.data_ptr(), B, 3, OH, OW, RH, RW, C.cast(m, _V) if m is not None else None,
                                                 C.cast(s, _V) if s is not None else None, int(swap_rb), int(chw),
                                                 int(dtype == torch.bfloat16), _nl.stream_ptr()), "zoo_serving_preproc")
    return out


def topk(x: torch.Tensor, k: int, softmax: bool = False):
    """Per-row top-k of a [B, N] tensor -> (indices int32 [B,k], values fp32 [B,k]), best first."""
    if x.dim() != 2:
        x = x.reshape(x.shape[0], -1)
    if not x.is_cuda or k > 8:
        v = torch.softmax(x.float(), -1) if softmax else x.float()
        t = v.topk(k, dim=-1)
        return t.indices.int(), t.values
    if x.dtype not in (torch.float32, torch.bfloat16):
        x = x.float()
    if x.stride(-1) != 1:
        x = x.contiguous()
    B, N = x.shape
    idx = torch.empty((B, k), dtype=torch.int32, device=x.device)
    val = torch.empty((B, k), dtype=torch.float32, device=x.device)
    _nl.check(_nl.cuda_lib().zoo_serving_topk(x.data_ptr(), x.stride(0), B, N, k, 0 if x.dtype == torch.float32 else 1,
                                              int(softmax), idx.data_ptr(), val.data_ptr(), _nl.stream_ptr()), "zoo_serving_topk")
    return idx, val


def format_topn(idx, val):
    """The reference's result string for one row: ``[[idx,prob],...]`` (PostProcessing.scala:83-92)."""
    return "[" + ",".join(f"[{int(i)},{float(v)}]" for i, v in zip(idx, val)) + "]"
