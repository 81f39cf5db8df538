"""InferenceModel: a thread-safe pool of weight-sharing execution slots.

Reference: pyzoo/zoo/pipeline/inference/inference_model.py:24-211 and
zoo/src/main/scala/.../pipeline/inference/InferenceModel.scala:30-842 -- a
`LinkedBlockingQueue` of `supportedConcurrentNum` model clones that share weights; `predict` takes
a clone, runs forward, offers it back; with auto-scaling, clones are made on demand.

Here a "clone" is an execution slot: its own CUDA stream plus (for static shapes) a CUDA graph of
the forward pass captured into slot-private input/output buffers.  Weights are shared by
construction (one module).  Several threads calling `predict` therefore overlap on the GPU the way
the reference's clones overlap on CPU cores.  Foreign formats are imported into torch graphs (Caffe prototxt /
caffemodel, TensorFlow frozen graphs); the reference's 8-bit engines (OpenVINO int8 IR, BigDL int8) are replaced by
`quantize()`: FP8 Linear layers on the tcgen05 kernel.
"""
from __future__ import annotations

import queue
import threading
import time

import numpy as np
import torch
import torch.nn as nn


class _Slot:
    def __init__(self, device, use_graph):
        self.stream = torch.cuda.Stream(device) if device.type == "cuda" else None
        self.use_graph = use_graph and device.type == "cuda"
        self.graphs = {}      # shape key -> (graph, static_inputs, static_output)
        self.seen = {}


class InferenceModel:
    def __init__(self, supported_concurrent_num=1, device=None, dtype=None, use_cuda_graph=True, channels_last=False, cast_weights=False):
        self.concurrent_num = int(supported_concurrent_num)
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.dtype = dtype                     # None = model's own precision; torch.bfloat16 = autocast
        self.use_cuda_graph = use_cuda_graph
        self.channels_last = channels_last
        # cast_weights: convert floating-point parameters and buffers to `dtype` ONCE at load time and feed inputs in that dtype,
        # instead of autocasting every call (autocast re-casts fp32 weights per forward; for inference the copy is pure overhead)
        self.cast_weights = bool(cast_weights) and dtype is not None
        self.model: nn.Module | None = None
        self._slots: queue.Queue | None = None
        self._made = 0
        self._lock = threading.Lock()
        self.summary = None
        self.records, self.batches, self.busy_s = 0, 0, 0.0

    # ---------------------------------------------------------------- loaders --
    def _install(self, model: nn.Module):
        model = model.to(self.device).eval()
        for p in model.parameters():
            p.requires_grad_(False)
        if self.channels_last == "auto":
            self.channels_last = any(isinstance(m, nn.Conv2d) for m in model.modules())
        if self.channels_last:
            model = model.to(memory_format=torch.channels_last)
        if self.cast_weights and self.device.type == "cuda":
            model = model.to(self.dtype)
        self.model = model
        n = self.concurrent_num if self.concurrent_num > 0 else 1
        self._slots = queue.Queue()
        self._made = 0
        for _ in range(n):
            self._slots.put(_Slot(self.device, self.use_cuda_graph))
            self._made += 1
        return self

    def load_torch(self, model):
        """A module object, or a path to a pickled module / TorchScript archive."""
        if isinstance(model, nn.Module):
            return self._install(model)
        try:
            return self._install(torch.jit.load(model, map_location="cpu"))
        except Exception:
            return self._install(torch.load(model, map_location="cpu", weights_only=False))

    load_torchscript = load_torch

    def load(self, model_path, weight_path=None):
        """A model written by `ZooModel.save_model` / `KerasNet.saveModel` (reference: load_bigdl)."""
        from ...models.common.zoo_model import ZooModel

        with open(model_path, "rb") as fh:
            head = fh.read(2)
        if head not in (b"PK", b"\x80\x02", b"\x80\x03", b"\x80\x04", b"\x80\x05"):      # BigDL's protobuf format
            from ...models.bigdl import load_bigdl

            return self._install(load_bigdl(model_path))
        try:
            m = ZooModel.load_model(model_path, weight_path)
        except Exception:
            m = torch.load(model_path, map_location="cpu", weights_only=False)
            if weight_path is not None:
                m.load_state_dict(torch.load(weight_path, map_location="cpu"))
        return self._install(m)

    load_bigdl = load

    def load_caffe(self, model_path, weight_path):
        """prototxt + caffemodel, converted to a torch graph (zoo.models.caffe; reference: doLoadCaffe)."""
        from ...models.caffe import load_caffe

        return self._install(load_caffe(model_path, weight_path))

    def load_tensorflow(self, model_path, model_type="frozenModel", intra_op_parallelism_threads=1, inter_op_parallelism_threads=1,
                        use_per_session_threads=True, inputs=None, outputs=None, tag=None, signature=None):
        """`model_type` "frozenModel": a frozen GraphDef (`.pb` + input / output names) or an export folder with graph_meta.json;
        "savedModel": a SavedModel directory (saved_model.pb + variables/), variables read from the bundle and frozen
        (zoo.tfpark.saved_model).  Both are imported into torch (zoo.tfpark.graphdef; reference: doLoadTF / doLoadTensorflow)."""
        import os

        from ..api.net.tfnet import TFNet

        kind = (model_type or "frozenModel").replace("_", "").lower()
        if kind == "savedmodel" or (os.path.isdir(model_path) and os.path.exists(os.path.join(model_path, "saved_model.pb"))):
            return self._install(TFNet.from_saved_model(model_path, tag=tag, signature=signature, inputs=inputs, outputs=outputs))
        if kind != "frozenmodel":
            raise ValueError(f"load_tensorflow: model_type {model_type!r} (expected 'frozenModel' or 'savedModel')")
        return self._install(TFNet(model_path, inputs, outputs))

    def load_tf(self, model_path, backend="tensorflow", intra_op_parallelism_threads=1, inter_op_parallelism_threads=1,
                use_per_session_threads=True, inputs=None, outputs=None, model_type=None, tag=None, signature=None):
        return self.load_tensorflow(model_path, model_type or "frozenModel", inputs=inputs, outputs=outputs, tag=tag, signature=signature)

    def load_openvino(self, model_path, weight_path=None, batch_size=0):
        """OpenVINO IR (`.xml` + `.bin`), the reference's CPU inference engine format (doLoadOpenVINO): the IR graph is imported as a
        torch module (`zoo.pipeline.inference.openvino_ir`, both the <blobs> generation the reference used and IR v10+); its 8-bit
        mode is `quantize(calibration_batch)` afterwards -- FP8 on the tensor cores."""
        from .openvino_ir import load_openvino_ir

        return self._install(load_openvino_ir(model_path, weight_path))

    def load_tf_as_openvino(self, *a, **kw):
        raise NotImplementedError("the *_as_openvino loaders run OpenVINO's Model Optimizer on a TensorFlow model; here: load_tf(...) followed by "
                                  "quantize(calibration_batch)")

    load_tf_object_detection_as_openvino = load_tf_image_classification_as_openvino = load_tf_as_calibrated_openvino = load_tf_as_openvino

    def quantize(self, calib_inputs, skip=()):
        """8-bit inference (reference: doLoadOpenVINOInt8 / BigDL int8 quantisation, OpenVinoInferenceSupportive.scala:35-57): every
        Linear / Dense layer of the loaded model becomes an FP8 (E4M3) layer on the tcgen05 kernel, with activation scales taken
        from `calib_inputs` (a tensor, a tuple of tensors, or a list of those).  See zoo.ops.quant."""
        from ...ops.quant import quantize_linear_layers

        if self.model is None:
            raise RuntimeError("load a model first")
        move = lambda t: t.to(self.device) if torch.is_tensor(t) else torch.as_tensor(np.asarray(t)).to(self.device)
        batches = calib_inputs if isinstance(calib_inputs, list) else [calib_inputs]
        batches = [tuple(move(t) for t in b) if isinstance(b, tuple) else move(b) for b in batches]
        quantize_linear_layers(self.model, batches, skip)
        for _ in range(self._slots.qsize()):          # captured graphs of the old layers are stale
            sl = self._slots.get()
            sl.graphs.clear(); sl.seen.clear()
            self._slots.put(sl)
        return self

    # ---------------------------------------------------------------- predict --
    def _take(self):
        if self.concurrent_num > 0:
            return self._slots.get()
        try:
            return self._slots.get_nowait()
        except queue.Empty:          # auto-scaling: make a slot on demand (InferenceModel.scala:784-803)
            with self._lock:
                self._made += 1
            return _Slot(self.device, self.use_cuda_graph)

    def _forward(self, xs):
        if self.cast_weights and self.device.type == "cuda":
            return self.model(*[x.to(self.dtype) if x.is_floating_point() else x for x in xs])
        if self.dtype is not None and self.device.type == "cuda":
            with torch.autocast("cuda", dtype=self.dtype):
                return self.model(*xs)
        return self.model(*xs)

    def _run_slot(self, slot, xs):
        if slot.stream is None:
            return self._forward(xs)
        key = tuple((tuple(x.shape), x.dtype) for x in xs)
        with torch.cuda.stream(slot.stream):
            slot.stream.wait_stream(torch.cuda.current_stream(self.device))
            g = slot.graphs.get(key)
            if g is None and slot.use_graph:
                slot.seen[key] = slot.seen.get(key, 0) + 1
                if slot.seen[key] == 3:     # shape is recurring: capture it
                    try:
                        static_in = [x.clone() for x in xs]
                        self._forward(static_in)
                        slot.stream.synchronize()
                        graph = torch.cuda.CUDAGraph()
                        # thread-local capture mode: other threads of the process (a serving engine's decode thread, other
                        # slots) keep making CUDA calls while this stream captures; the default global mode makes those fail
                        with torch.cuda.graph(graph, stream=slot.stream, capture_error_mode="thread_local"):
                            static_out = self._forward(static_in)
                        g = slot.graphs[key] = (graph, static_in, static_out)
                    except Exception:
                        slot.use_graph = False
                        slot.stream.synchronize()
            if g is not None:
                graph, static_in, static_out = g
                for d, s in zip(static_in, xs):
                    d.copy_(s, non_blocking=True)
                graph.replay()
                out = static_out.clone() if torch.is_tensor(static_out) else tuple(o.clone() for o in static_out)
            else:
                out = self._forward(xs)
        torch.cuda.current_stream(self.device).wait_stream(slot.stream)
        return out

    @torch.no_grad()
    def predict_tensor(self, *xs):
        """Device tensors in, device tensor(s) out (the serving engine's path)."""
        if self.model is None:
            raise RuntimeError("no model loaded")
        slot = self._take()
        t0 = time.perf_counter()
        try:
            xs = [x.to(self.device, non_blocking=True) for x in xs]
            if self.channels_last:
                xs = [x.contiguous(memory_format=torch.channels_last) if x.dim() == 4 else x for x in xs]
            return self._run_slot(slot, xs)
        finally:
            self._slots.put(slot)
            with self._lock:
                self.batches += 1
                self.records += int(xs[0].shape[0]) if xs and xs[0].dim() else 1
                self.busy_s += time.perf_counter() - t0

    def predict(self, inputs):
        """ndarray | tensor | list of them (multi-input) -> ndarray (or list of ndarrays)."""
        multi = isinstance(inputs, (list, tuple)) and len(inputs) > 0 and all(isinstance(i, np.ndarray) or torch.is_tensor(i) for i in inputs)
        xs = list(inputs) if multi else [inputs]
        ts = [x if torch.is_tensor(x) else torch.as_tensor(np.asarray(x)) for x in xs]
        ts = [t.float() if t.dtype == torch.float64 else t for t in ts]
        out = self.predict_tensor(*ts)
        if torch.is_tensor(out):
            return out.float().cpu().numpy()
        return [o.float().cpu().numpy() for o in out]

    do_predict = predict

    def release(self):
        self.model, self._slots = None, None

    def __repr__(self):
        return f"InferenceModel(concurrent_num={self.concurrent_num}, slots={self._made}, device={self.device}, dtype={self.dtype})"
