"""InferenceModel: a thread-safe pool of weight-sharing execution slots.

Reference: pyzoo/zoo/pipeline/inference/inference_model.py:24-211 and
zoo/src/main/scala/.../pipeline/inference/InferenceModel.scala:30-842 -- a
`LinkedBlockingQueue` of `supportedConcurrentNum` model clones that share weights; `predict` takes
a clone, runs forward, offers it back; with auto-scaling, clones are made on demand.

Here a "clone" is an execution slot: its own CUDA stream plus (for static shapes) a CUDA graph of
the forward pass captured into slot-private input/output buffers.  Weights are shared by
construction (one module).  Several threads calling `predict` therefore overlap on the GPU the way
the reference's clones overlap on CPU cores.  Loaders for engines that do not exist on this box
(Caffe, OpenVINO, TensorFlow) raise `NotImplementedError` naming the supported route
(TorchScript / a pickled zoo model / a module object).
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
    def __init__(self, supported_concurrent_num=1, device=None, dtype=None, use_cuda_graph=True, channels_last=False):
        self.concurrent_num = int(supported_concurrent_num)
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.dtype = dtype                     # None = model's own precision; torch.bfloat16 = autocast
        self.use_cuda_graph = use_cuda_graph
        self.channels_last = channels_last
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
        if self.channels_last:
            model = model.to(memory_format=torch.channels_last)
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

        try:
            m = ZooModel.load_model(model_path, weight_path)
        except Exception:
            m = torch.load(model_path, map_location="cpu", weights_only=False)
            if weight_path is not None:
                m.load_state_dict(torch.load(weight_path, map_location="cpu"))
        return self._install(m)

    load_bigdl = load

    def _unavailable(self, what):
        raise NotImplementedError(f"{what} needs an engine that is not part of zoo-b200 (no TensorFlow/Caffe/OpenVINO runtime); "
                                  "export the model to TorchScript and use load_torch(path)")

    def load_caffe(self, model_path, weight_path):
        self._unavailable("load_caffe")

    def load_openvino(self, model_path, weight_path, batch_size=0):
        self._unavailable("load_openvino")

    def load_tensorflow(self, *a, **kw):
        self._unavailable("load_tensorflow")

    load_tf = load_tf_object_detection_as_openvino = load_tf_image_classification_as_openvino = load_tensorflow
    load_tf_as_calibrated_openvino = load_tensorflow

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
                        with torch.cuda.graph(graph, stream=slot.stream):
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
