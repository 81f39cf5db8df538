"""Cluster Serving client: `InputQueue.enqueue_image(uri, img)` / `OutputQueue.query(uri)` / `dequeue()`.

Reference: pyzoo/zoo/serving/client.py:25-151.  Same behaviour -- images are resized to the configured
shape, JPEG-encoded, base64-encoded and `XADD`ed to `image_stream` as {uri, image}; pushes are refused
while the broker's memory use is above 60 % of `maxmemory`; results are the hashes `result:<uri>` with
a `value` field.  Differences: the connection is either the in-process broker of a `ClusterServing`
object (no socket) or a RESP socket to `data.src`; and `enqueue_tensor` accepts any ndarray (raw
little-endian bytes + shape + dtype), which is how NCF and other non-image models are served.
"""
from __future__ import annotations

import base64
import os
import time

import numpy as np
import yaml

from ..native.rt import RespClient, RespError

INPUT_STREAM = "image_stream"
GROUP = "serving"


def load_config(path="config.yaml"):
    if isinstance(path, dict):
        cfg = path
    elif path and os.path.exists(path):
        with open(path) as f:
            cfg = yaml.safe_load(f) or {}
    else:
        cfg = {}
    for sec in ("model", "data", "params", "redis"):
        if not isinstance(cfg.get(sec), dict):
            cfg[sec] = {}
    return cfg


class API:
    """Connection + config shared by the two queues (client.py:25-59)."""

    def __init__(self, config="config.yaml", broker=None, host=None, port=None):
        cfg = load_config(config)
        src = cfg["data"].get("src") or "localhost:6379"
        h, _, p = str(src).partition(":")
        self.host, self.port = host or h or "localhost", int(port or p or 6379)
        self.db = broker if broker is not None else RespClient("127.0.0.1" if self.host == "localhost" else self.host, self.port)
        try:
            self.db.xgroup_create(INPUT_STREAM, GROUP, id="$", mkstream=True)
        except RespError:
            pass  # the group exists
        shape = cfg["data"].get("image_shape") or "3,224,224"
        self.data_shape = [int(v) for v in str(shape).replace(" ", "").split(",")]


class InputQueue(API):
    def __init__(self, config="config.yaml", broker=None, host=None, port=None):
        super().__init__(config, broker, host, port)
        self.c = self.h = self.w = None
        self.input_threshold = 0.6
        self.interval_if_error = 1
        self.data_shape_check()

    def data_shape_check(self):
        for num in self.data_shape:
            if num <= 0:
                raise Exception(f"Your image shape config is invalid, your config shape is {self.data_shape}; no negative value is allowed.")
            if 0 < num < 5:
                self.c = num
                continue
            if not self.h:
                self.h = num
            else:
                self.w = num

    def _guarded_xadd(self, fields):
        """Refuse while used_memory >= 60 % of maxmemory (client.py:108-121).  Returns True when written."""
        inf = self.db.info()
        if inf.get("maxmemory", 0) and inf["used_memory"] >= inf["maxmemory"] * self.input_threshold:
            print("Redis queue is full, please wait for inference or delete the unprocessed records.")
            time.sleep(self.interval_if_error)
            return False
        try:
            self.db.xadd(INPUT_STREAM, fields)
            return True
        except RespError as e:
            print(e, "Redis memory is full, please dequeue or delete.")
            time.sleep(self.interval_if_error)
            return False

    def enqueue_image(self, uri, img, resize=True):
        import cv2

        if isinstance(img, str):
            path = img
            img = cv2.imread(path)
            if img is None:
                print("You have pushed an image with path: ", path, "the path is invalid, skipped.")
                return False
        if resize:   # force one shape so records of a batch agree (client.py:100-102)
            img = cv2.resize(img, (self.w, self.h))
        ok, data = cv2.imencode(".jpg", img)
        if not ok:
            raise ValueError("JPEG encoding failed")
        return self._guarded_xadd({"uri": uri, "image": self.base64_encode_image(data)})

    def enqueue_tensor(self, uri, arr):
        a = np.ascontiguousarray(arr)
        return self._guarded_xadd({"uri": uri, "tensor": a.tobytes(), "shape": ",".join(map(str, a.shape)), "dtype": a.dtype.str})

    def enqueue_tensors(self, uris, arr):
        """Bulk form of `enqueue_tensor`: one record per row of `arr`.  With an in-process broker this is a single native
        call; over a socket it falls back to one XADD per record.  Returns the number of records written."""
        a = np.ascontiguousarray(arr)
        if hasattr(self.db, "xadd_many"):
            inf = self.db.info()
            if inf.get("maxmemory", 0) and inf["used_memory"] >= inf["maxmemory"] * self.input_threshold:
                print("Redis queue is full, please wait for inference or delete the unprocessed records.")
                time.sleep(self.interval_if_error)
                return 0
            return self.db.xadd_many(INPUT_STREAM, list(uris), "tensor", a, {"shape": ",".join(map(str, a.shape[1:])), "dtype": a.dtype.str})
        return sum(bool(self.enqueue_tensor(u, a[k])) for k, u in enumerate(uris))

    @staticmethod
    def base64_encode_image(img):
        return base64.b64encode(img).decode("utf-8")


class OutputQueue(API):
    def dequeue(self):
        decoded = {}
        for res in self.db.keys("result:*"):
            key = res.decode("utf-8")
            res_dict = self.db.hgetall(key)
            if b"value" in res_dict:
                decoded[key.split(":", 1)[1]] = res_dict[b"value"].decode("utf-8")
            self.db.delete(key)
        return decoded

    def query(self, uri):
        res_dict = self.db.hgetall("result:" + uri)
        if not res_dict:
            return "{}"
        return res_dict[b"value"].decode("utf-8")
