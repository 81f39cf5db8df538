"""Cluster Serving engine: broker -> decode -> GPU preprocess -> model -> top-N -> broker.

Reference: zoo/src/main/scala/.../serving/ClusterServing.scala:44-352 (+ utils/ClusterServingHelper.scala,
ClusterServingManager.scala, PreProcessing.scala, PostProcessing.scala).  The reference is a Spark
Structured-Streaming job: every micro-batch is read from Redis, decoded with OpenCV, copied into a
`[batch,C,H,W]` tensor, run through an `InferenceModel` clone, post-processed per row and written back
as `result:<uri>` hashes; a `running` file stops it; input is trimmed when Redis passes 48 % of its
memory; throughput goes to TensorBoard (`Serving Throughput`, `Total Records Number`).

Here one process per GPU runs a three-stage pipeline over the in-process broker
(`csrc/runtime/broker.cpp`): (1) `XREADGROUP` whatever is queued, up to `batch_size` (continuous
batching: a partial batch is served at once, never waited on), JPEG-decode on a thread pool;
(2) one H2D copy of the packed pixels, kernel K13 (resize / normalise / layout), the model forward on an
`InferenceModel` slot (CUDA-graph replay for recurring batch shapes), kernel K14 (softmax + top-N);
(3) format + `HSET` on a writer thread.  Config keys, stream / result names, the `filter: topN(k)`
syntax, the stop file and the two memory thresholds are the reference's.
"""
from __future__ import annotations

import base64
import logging
import os
import queue
import re
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from ..native.rt import Broker, RespClient, RespError
from ..pipeline.inference import InferenceModel
from .client import GROUP, INPUT_STREAM, load_config

log = logging.getLogger("zoo.serving")


def _mem_bytes(s, default):
    if s in (None, ""):
        return default
    m = re.fullmatch(r"\s*(\d+)\s*([kmgKMG]?)[bB]?\s*", str(s))
    if not m:
        raise ValueError(f"bad memory size {s!r}")
    return int(m.group(1)) << {"": 0, "k": 10, "m": 20, "g": 30}[m.group(2).lower()]


class ServingConfig:
    """config.yaml -> typed settings (ClusterServingHelper.scala:101-169): defaults batch_size 4,
    image_shape 3,224,224, filter topN(1), maxmem 4g, src localhost:6379."""

    def __init__(self, config="config.yaml", **overrides):
        c = load_config(config)
        m, d, p, r = c["model"], c["data"], c["params"], c["redis"]
        self.model_path = overrides.get("model_path", m.get("path"))
        self.model_inputs, self.model_outputs = m.get("inputs"), m.get("outputs")
        src = str(d.get("src") or "localhost:6379")
        self.host, _, port = src.partition(":")
        self.port = int(port or 6379)
        self.image_shape = [int(v) for v in str(d.get("image_shape") or "3,224,224").replace(" ", "").split(",")]
        self.filter = "topN(1)" if d.get("filter") is None else str(d.get("filter"))   # blank key = default; "" / none = raw tensor
        self.batch_size = int(p.get("batch_size") or 4)
        self.performance_mode = str(p.get("performance_mode") or "OFF").upper() in ("ON", "TRUE", "1")
        self.maxmem = _mem_bytes(r.get("maxmem"), 4 << 30)
        self.mean = d.get("mean") or [123.68, 116.779, 103.939]
        self.std = d.get("std") or [58.393, 57.12, 57.375]
        self.log_dir = c.get("log_dir") or "."
        self.dtype = str(p.get("dtype") or "bf16")
        for k, v in overrides.items():
            setattr(self, k, v)
        chans = [v for v in self.image_shape if 0 < v < 5]
        sizes = [v for v in self.image_shape if v >= 5]
        self.C = chans[0] if chans else 3
        self.H, self.W = (sizes + sizes)[:2] if sizes else (224, 224)
        self.chw = self.image_shape.index(self.C) == 0 if chans else True
        self.top_n = self.parse_filter(self.filter)

    @staticmethod
    def parse_filter(f):
        """`topN(k)` -> k; empty / anything else -> None = return the raw tensor (PostProcessing.scala:96-116)."""
        m = re.fullmatch(r"\s*topN\((\d+)\)\s*", f or "")
        if m:
            return int(m.group(1))
        if f and f.strip() and f.strip().lower() != "none":
            raise ValueError(f"Filter string {f!r} is invalid: only topN(k) is supported")
        return None


class ClusterServing:
    def __init__(self, config="config.yaml", model=None, broker=None, device=None, serve_socket=False, consumer=None,
                 decode_threads=None, **overrides):
        self.conf = config if isinstance(config, ServingConfig) else ServingConfig(config, **overrides)
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.owns_broker = broker is None
        self.db = broker if broker is not None else Broker(self.conf.maxmem)
        if serve_socket and isinstance(self.db, Broker):
            self.port = self.db.serve("127.0.0.1" if self.conf.host == "localhost" else self.conf.host, self.conf.port)
        try:
            self.db.xgroup_create(INPUT_STREAM, GROUP, id="0", mkstream=True)
        except RespError:
            pass
        # params.dtype: bf16 | fp16 (autocast) | fp32 | fp8 (Linear / Dense layers on the FP8 tensor-core path, calibrated on the first
        # batch that is served -- the 8-bit route of the reference's OpenVINO int8 serving; the rest of the model runs in bf16)
        dt = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp8": torch.bfloat16}.get(self.conf.dtype) if self.device.type == "cuda" else None
        self._quantize_pending = self.conf.dtype == "fp8"
        # image models: NHWC (channels_last) activations and weights cast once to the compute dtype (no per-call autocast)
        self.model = InferenceModel(1, self.device, dtype=dt, channels_last="auto", cast_weights=dt is not None and self.conf.dtype != "fp8")
        if model is not None:
            self.model.load_torch(model)
        elif self.conf.model_path:
            self._load_by_extension(self.conf.model_path)
        else:
            raise ValueError("model.path is not set in the config and no model object was given")
        self.consumer = consumer or f"c{os.getpid()}"
        self.pool = ThreadPoolExecutor(decode_threads or min(16, os.cpu_count() or 4))
        self._stop = threading.Event()
        self._threads = []
        self._out_q = queue.Queue(maxsize=8)
        self.total_records = 0
        self.stage_s = {"read": 0.0, "b64": 0.0, "decode+preproc": 0.0, "model+post": 0.0, "batches": 0}   # host time per stage
        self.latencies = []           # per-batch (n, seconds in engine)
        self.summary = None
        self.stop_file = os.path.join(self.conf.log_dir, "running")
        self._t_start = None

    def _load_by_extension(self, path):
        """Model type from the file name (ClusterServingHelper.scala:350-448)."""
        if os.path.isdir(path):
            files = sorted(os.listdir(path))
            pick = [f for f in files if f.endswith((".pt", ".pth", ".model", ".bigdl", ".zoo"))]
            if not pick:
                raise FileNotFoundError(f"no supported model file under {path} (.pt/.pth TorchScript or pickled module, .model zoo model)")
            path = os.path.join(path, pick[0])
        if path.endswith((".pt", ".pth")):
            self.model.load_torch(path)
        elif path.endswith((".caffemodel", ".prototxt")):
            stem = path.rsplit(".", 1)[0]
            proto, weights = stem + ".prototxt", stem + ".caffemodel"
            self.model.load_caffe(proto, weights if os.path.exists(weights) else None)
        elif path.endswith(".pb"):
            folder = os.path.dirname(path)
            if os.path.exists(os.path.join(folder, "graph_meta.json")):
                self.model.load_tensorflow(folder)
            else:       # config.yaml: model.inputs / model.outputs name the tensors of a bare frozen graph
                split = lambda v: [t.strip() for t in str(v).split(",") if t.strip()] if v else None
                self.model.load_tensorflow(path, inputs=split(getattr(self.conf, "model_inputs", None)), outputs=split(getattr(self.conf,
                        "model_outputs", None)))
        elif path.endswith((".xml", ".bin")):
            self.model.load_openvino(path, None)
        else:
            self.model.load(path)

    # ------------------------------------------------------------------ stages --
    def _decode_one(self, fields):
        if "tensor" in fields:
            shape = [int(v) for v in fields["shape"].decode().split(",") if v]
            return np.frombuffer(fields["tensor"], dtype=np.dtype(fields["dtype"].decode())).reshape(shape)
        import cv2

        raw = base64.b64decode(fields["image"])
        img = cv2.imdecode(np.frombuffer(raw, np.uint8), cv2.IMREAD_COLOR if self.conf.C == 3 else cv2.IMREAD_GRAYSCALE)
        if img is None:
            raise ValueError("image could not be decoded")
        return img

    def _read_batch(self, block_ms):
        r = self.db.xreadgroup(GROUP, self.consumer, {INPUT_STREAM: ">"}, count=self.conf.batch_size, block=block_ms, noack=True)
        return r[0][1] if r else []

    def _memory_guard(self):
        """Trim the input stream to half when the broker is past 48 % of its memory (ClusterServing.scala:134-140)."""
        inf = self.db.info()
        if inf.get("maxmemory", 0) and inf["used_memory"] >= 0.48 * inf["maxmemory"]:
            n = self.db.xlen(INPUT_STREAM)
            self.db.xtrim(INPUT_STREAM, n // 2)
            log.warning("broker memory above 48%%: trimmed %s to %d entries", INPUT_STREAM, n // 2)

    def _predict(self, x):
        """Model forward for one batch; the first batch also calibrates the FP8 layers when params.dtype is fp8."""
        if self._quantize_pending:
            self._quantize_pending = False
            self.model.quantize(x)
        return self.model.predict_tensor(x)

    def _jpeg_decoder(self):
        """nvJPEG batched decoder on the engine's GPU (None on CPU, without libnvjpeg, or with ZOO_SERVING_NVJPEG=0)."""
        if not hasattr(self, "_jpeg"):
            self._jpeg = None
            if self.device.type == "cuda" and os.environ.get("ZOO_SERVING_NVJPEG", "1") != "0":
                from ..ops.serving_ops import JpegBatchDecoder

                if JpegBatchDecoder.available():
                    try:
                        self._jpeg = JpegBatchDecoder(self.device, max_batch=max(512, self.conf.batch_size))
                    except Exception as e:
                        log.warning("nvJPEG decoder unavailable (%r); decoding on host threads", e)
        return self._jpeg

    def _prep_jpegs(self, fields):
        """base64 (native threads) -> nvJPEG batched decode -> K13, all enqueued on the CURRENT stream; returns the model input."""
        from ..native.rt import b64_decode_batch
        from ..ops.serving_ops import preprocess_jpegs

        c = self.conf
        t0 = time.perf_counter()
        self._b64_buf, offs, lens = b64_decode_batch([f["image"] for f in fields], getattr(self, "_b64_buf", None),
                                                     threads=min(8, os.cpu_count() or 4))
        t1 = time.perf_counter()
        dtype = self.model.dtype or torch.float32
        # NHWC output when the model runs channels_last: K13 writes the layout the convolutions want, no transpose pass
        nhwc = bool(self.model.channels_last) and c.chw
        x = preprocess_jpegs((self._b64_buf, offs, lens), (c.H, c.W), self._jpeg, None, c.mean[:3], c.std[:3], swap_rb=True,
                             chw=not nhwc and c.chw, dtype=dtype)
        if nhwc:
            x = x.permute(0, 3, 1, 2)           # logical NCHW over channels_last memory
        st = self.stage_s
        st["b64"] += t1 - t0; st["decode+preproc"] += time.perf_counter() - t1
        return x

    def _infer_jpegs(self, fields):
        """JPEG records of one batch: base64 -> nvJPEG batched decode on the GPU -> K13 -> model -> K14; pixels never touch the host.
        A batch nvJPEG cannot decode (progressive / CMYK / corrupt records) is retried on the host path."""
        try:
            x = self._prep_jpegs(fields)
        except Exception as e:
            log.warning("nvJPEG batch decode failed (%r); decoding this batch on the host", e)
            return self._infer(list(self.pool.map(self._decode_one, fields)))
        t2 = time.perf_counter()
        res = self._finish(self._predict(x))
        self.stage_s["model+post"] += time.perf_counter() - t2
        self.stage_s["batches"] += 1
        return res

    def _prep_loop(self):
        """Pipelined image mode: this thread reads, base64-decodes and enqueues decode + K13 on its own CUDA stream while the engine
        thread runs the model of the previous batch; batches are handed over with a CUDA event (no host synchronisation)."""
        self._bind_device()
        stream = torch.cuda.Stream(self.device)
        while not self._stop.is_set():
            tr = time.perf_counter()
            entries = self._read_batch(5)
            if not entries:
                continue
            self.stage_s["read"] += time.perf_counter() - tr
            t0 = time.perf_counter()
            uris = [f["uri"].decode() for _, f in entries]
            fields = [f for _, f in entries]
            item = None
            if all("image" in f for f in fields):
                try:
                    with torch.cuda.stream(stream):
                        x = self._prep_jpegs(fields)
                        ev = torch.cuda.Event()
                        ev.record(stream)
                    item = (uris, x, ev, None, t0)
                except Exception as e:
                    log.warning("nvJPEG batch decode failed (%r); decoding this batch on the host", e)
            if item is None:
                item = (uris, None, None, fields, t0)
            while not self._stop.is_set():
                try:
                    self._prep_q.put(item, timeout=0.2)
                    break
                except queue.Full:
                    continue

    def _engine_from_prep(self):
        """Engine side of the pipelined image mode."""
        cur = torch.cuda.current_stream(self.device)
        while not self._stop.is_set():
            try:
                uris, x, ev, fields, t0 = self._prep_q.get(timeout=0.2)
            except queue.Empty:
                continue
            try:
                if x is not None:
                    cur.wait_event(ev)
                    x.record_stream(cur)
                    t2 = time.perf_counter()
                    kind, payload = self._finish(self._predict(x))
                    self.stage_s["model+post"] += time.perf_counter() - t2
                    self.stage_s["batches"] += 1
                else:
                    kind, payload = self._infer(list(self.pool.map(self._decode_one, fields)))
            except Exception as e:
                log.exception("batch failed: %r", e)
                self._out_q.put((uris, "error", repr(e), t0))
                continue
            self._out_q.put((uris, kind, payload, t0))
            if self.total_records % 2048 < len(uris):
                self._memory_guard()

    def _infer(self, items):
        """items: list of decoded arrays -> (kind, payload) for the writer."""
        c = self.conf
        first = items[0]
        is_image = first.dtype == np.uint8 and first.ndim in (2, 3)
        if is_image:
            from ..ops.serving_ops import preprocess_images

            dtype = self.model.dtype or torch.float32
            x = preprocess_images(items, (c.H, c.W), None, c.mean[:c.C], c.std[:c.C], swap_rb=(c.C == 3), chw=c.chw,
                                  dtype=dtype if self.device.type == "cuda" else torch.float32, device=self.device)
            out = self._predict(x)
        else:
            xs = torch.from_numpy(np.stack(items))
            out = self._predict(xs)
        return self._finish(out)

    def _finish(self, out):
        """Model output -> (kind, payload) for the writer: per-row top-N on the GPU (K14) or the raw tensor."""
        c = self.conf
        if isinstance(out, (tuple, list)):
            out = out[0]
        out = out.reshape(out.shape[0], -1)
        if c.top_n is not None:
            from ..ops.serving_ops import topk

            idx, val = topk(out, min(c.top_n, out.shape[1], 8), softmax=False)
            return "topn", (idx.cpu().numpy(), val.cpu().numpy())
        return "raw", out.float().cpu().numpy()

    def _packed_batch(self):
        """Tensor records through the broker's batch path: one native call returns the batch as one contiguous array."""
        if self._pack_buf is None:
            self._pack_buf = np.empty(max(16 << 20, self.conf.batch_size * 4096), dtype=np.uint8)
            self._uri_buf = np.empty(max(1 << 20, self.conf.batch_size * 96), dtype=np.uint8)
        n, data, uris = self.db.read_pack(INPUT_STREAM, GROUP, self.conf.batch_size, "tensor", 5, self._pack_buf, self._uri_buf,
                                          raw_uris=True)          # the uri blob goes to the result writer undecoded
        if n == 0:
            return None
        shape, dtype = self._tensor_meta
        return uris, data.view(dtype).reshape(n, *shape)

    def _bind_device(self):
        """The current CUDA device is per THREAD: a worker thread starts on device 0, and a kernel launched from it through the
        native library would run there while its pointers live on this rank's device."""
        if self.device.type == "cuda" and self.device.index is not None:
            torch.cuda.set_device(self.device)

    def _engine_loop(self):
        self._bind_device()
        idle = 0
        self._pack_buf, self._tensor_meta = None, None
        while not self._stop.is_set():
            if self._tensor_meta is not None and hasattr(self.db, "read_pack"):
                got = self._packed_batch()
                if got is None:
                    continue
                uris, batch = got
                t0 = time.perf_counter()
                try:
                    out = self._predict(torch.from_numpy(batch))
                    out = (out[0] if isinstance(out, (tuple, list)) else out)
                    out = out.reshape(out.shape[0], -1)
                    if self.conf.top_n is not None:
                        from ..ops.serving_ops import topk

                        idx, val = topk(out, min(self.conf.top_n, out.shape[1], 8), softmax=False)
                        self._out_q.put((uris, "topn", (idx.cpu().numpy(), val.cpu().numpy()), t0))
                    else:
                        self._out_q.put((uris.decode().split("\n")[:-1], "raw", out.float().cpu().numpy(), t0))
                except Exception as e:
                    log.exception("batch failed: %r", e)
                    self._out_q.put((uris.decode().split("\n")[:-1], "error", repr(e), t0))
                continue
            tr = time.perf_counter()
            entries = self._read_batch(5)
            if entries:
                self.stage_s["read"] += time.perf_counter() - tr
            if not entries:
                idle += 1
                if idle % 200 == 0 and not os.path.exists(self.stop_file) and getattr(self, "_watch_file", False):
                    self._stop.set()          # "running" file removed: stop (ClusterServingManager.scala:32-53)
                continue
            t0 = time.perf_counter()
            uris = [f["uri"].decode() for _, f in entries]
            f0 = entries[0][1]
            if "tensor" in f0 and self._tensor_meta is None and hasattr(self.db, "read_pack"):
                # tensor records: from the next batch on use the packed batch path (homogeneous shape / dtype)
                self._tensor_meta = ([int(v) for v in f0["shape"].decode().split(",") if v], np.dtype(f0["dtype"].decode()))
            try:
                if "image" in f0 and self._jpeg_decoder() is not None and self.conf.C == 3:
                    kind, payload = self._infer_jpegs([f for _, f in entries])
                    if os.environ.get("ZOO_SERVING_PIPELINE", "1") != "0":
                        # image stream on a GPU with nvJPEG: from here on a second thread prepares batch i+1 while this one runs batch i
                        self._out_q.put((uris, kind, payload, t0))
                        self._prep_q = queue.Queue(maxsize=2)
                        t = threading.Thread(target=self._prep_loop, daemon=True, name="zoo-serving-prep")
                        t.start()
                        self._threads.append(t)
                        return self._engine_from_prep()
                else:
                    items = list(self.pool.map(self._decode_one, [f for _, f in entries]))
                    kind, payload = self._infer(items)
            except Exception as e:  # a bad record must not stop the service
                log.exception("batch failed: %r", e)
                self._out_q.put((uris, "error", repr(e), t0))
                continue
            self._out_q.put((uris, kind, payload, t0))
            if self.total_records % 2048 < len(uris):
                self._memory_guard()

    def _writer_loop(self):
        from ..ops.serving_ops import format_topn

        self._bind_device()
        while True:
            item = self._out_q.get()
            if item is None:
                return
            uris, kind, payload, t0 = item
            packed = isinstance(uris, (bytes, bytearray))        # the undecoded uri blob of the packed batch path
            count = uris.count(b"\n") if packed else len(uris)
            values = None
            if packed:
                pass                                             # formatted natively by hset_topn
            elif kind == "topn":
                if payload[0].shape[1] == 1:
                    values = [f"[[{i},{v}]]" for i, v in zip(payload[0][:, 0].tolist(), payload[1][:, 0].tolist())]
                else:
                    values = [format_topn(payload[0][k], payload[1][k]) for k in range(len(uris))]
            elif kind == "raw":
                values = [np.array2string(payload[k], separator=",", threshold=1 << 20, max_line_width=1 << 20) for k in range(len(uris))]
            else:
                values = ["NaN"] * len(uris)
            while True:          # retry forever on a full broker, every 3 s (ClusterServing.scala:276-310)
                try:
                    if packed:
                        self.db.hset_topn("result:", "value", bytes(uris), payload[0], payload[1])
                    elif hasattr(self.db, "hset_many"):
                        self.db.hset_many("result:", "value", uris, values)
                    else:
                        for uri, value in zip(uris, values):
                            self.db.hset("result:" + uri, "value", value)
                    break
                except RespError as e:
                    log.warning("result write failed (%s); retrying in 3 s", e)
                    if self._stop.wait(3):
                        break
            dt = time.perf_counter() - t0
            self.total_records += count
            self.latencies.append((count, dt))
            if self.summary is not None:
                el = max(time.perf_counter() - self._t_start, 1e-9)
                self.summary.add_scalar("Serving Throughput", self.total_records / el, self.total_records)
                self.summary.add_scalar("Total Records Number", self.total_records, self.total_records)

    # ----------------------------------------------------------------- control --
    def start(self, watch_stop_file=False, tensorboard=False):
        self._stop.clear()
        self._watch_file = watch_stop_file
        if watch_stop_file:
            os.makedirs(self.conf.log_dir, exist_ok=True)
            open(self.stop_file, "w").close()
        if tensorboard:
            from ..tensorboard.writer import InferenceSummary

            self.summary = InferenceSummary(self.conf.log_dir, "cluster-serving")
        self._t_start = time.perf_counter()
        self._threads = [threading.Thread(target=self._engine_loop, daemon=True, name="zoo-serving-engine"),
                         threading.Thread(target=self._writer_loop, daemon=True, name="zoo-serving-writer")]
        for t in self._threads:
            t.start()
        return self

    def stop(self):
        self._stop.set()
        if self._threads:
            self._threads[0].join(10)
            self._out_q.put(None)
            self._threads[1].join(10)
            self._threads = []
        if self.summary is not None:
            self.summary.close()

    def close(self):
        self.stop()
        self.pool.shutdown(wait=False)
        if self.owns_broker and isinstance(self.db, Broker):
            self.db.stop_server()
            self.db.close()

    def wait_until(self, n_records, timeout=60.0):
        t0 = time.time()
        while self.total_records < n_records and time.time() - t0 < timeout:
            time.sleep(0.001)
        return self.total_records >= n_records

    def stats(self):
        lat = np.array([d for _, d in self.latencies]) if self.latencies else np.zeros(1)
        el = max(time.perf_counter() - (self._t_start or time.perf_counter()), 1e-9)
        return {"records": self.total_records, "throughput": self.total_records / el,
                "p50_ms": float(np.percentile(lat, 50) * 1e3), "p99_ms": float(np.percentile(lat, 99) * 1e3)}

    def serve_forever(self):
        self.start(watch_stop_file=True, tensorboard=True)
        try:
            while not self._stop.wait(1.0):
                if not os.path.exists(self.stop_file):
                    break
        finally:
            self.close()


def main(argv=None):
    """`python -m zoo.serving.server [-c config.yaml]` -- what `cluster-serving-start` runs."""
    import argparse

    ap = argparse.ArgumentParser()
    ap.add_argument("-c", "--config", default="config.yaml")
    ap.add_argument("--port", type=int, default=None, help="RESP listener port (overrides data.src of the config; one per GPU process)")
    a = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, filename="cluster-serving.log")
    conf = ServingConfig(a.config)
    if a.port is not None:
        conf.port = a.port
    ClusterServing(conf, serve_socket=True).serve_forever()


if __name__ == "__main__":
    main()
