"""Cluster Serving end to end on CPU: enqueue -> engine -> query/dequeue, config parsing, filters,
back-pressure, InferenceModel pool semantics (reference: zoo/src/test/scala/.../serving/*Spec.scala and the
offline-benchmark harness)."""
import threading
import time

import numpy as np
import pytest
import torch
import torch.nn as nn


class TinyCls(nn.Module):
    def __init__(self, n=10):
        super().__init__()
        self.conv = nn.Conv2d(3, 4, 3, padding=1)
        self.fc = nn.Linear(4, n)

    def forward(self, x):
        return torch.softmax(self.fc(torch.relu(self.conv(x.float())).mean((2, 3))), -1)


def test_config_and_filter_parsing(tmp_path):
    from zoo.serving.server import ServingConfig

    p = tmp_path / "config.yaml"
    p.write_text("model:\n  path: /m\ndata:\n  src: 10.0.0.1:7000\n  image_shape: 3, 32,48\n  filter: topN(3)\n"
                 "params:\n  batch_size: 16\nredis:\n  maxmem: 2g\n")
    c = ServingConfig(str(p))
    assert (c.host, c.port, c.C, c.H, c.W, c.top_n, c.batch_size, c.maxmem, c.chw) == ("10.0.0.1", 7000, 3, 32, 48, 3, 16, 2 << 30, True)
    d = ServingConfig({})
    assert (d.port, d.batch_size, d.top_n, d.image_shape) == (6379, 4, 1, [3, 224, 224])
    with pytest.raises(ValueError):
        ServingConfig({"data": {"filter": "bogus(1)"}})
    assert ServingConfig({"data": {"image_shape": "224,224,3"}}).chw is False


def test_inference_model_pool():
    from zoo.pipeline.inference import InferenceModel

    m = InferenceModel(supported_concurrent_num=2, device="cpu").load_torch(TinyCls())
    x = np.random.rand(5, 3, 8, 8).astype(np.float32)
    ref = m.predict(x)
    assert ref.shape == (5, 10)
    outs = [None] * 8

    def work(i):
        outs[i] = m.predict(x)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(8)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert all(np.allclose(o, ref) for o in outs) and m.records == 45
    auto = InferenceModel(supported_concurrent_num=0, device="cpu").load_torch(TinyCls())
    assert auto.predict(x).shape == (5, 10)
    with pytest.raises(FileNotFoundError):
        m.load_openvino("a.xml", "a.bin")                   # an IR importer now (tests/test_cpu_openvino_ir.py): a missing file is just missing

    class Two(nn.Module):
        def forward(self, a, b):
            return a + b.sum(1, keepdim=True)

    two = InferenceModel(device="cpu").load_torch(Two())
    np.testing.assert_allclose(two.predict([np.ones((2, 3), np.float32), np.ones((2, 4), np.float32)]), np.full((2, 3), 5.0))


def test_serving_end_to_end_images_and_tensors(tmp_path):
    from zoo.serving import ClusterServing, InputQueue, OutputQueue

    torch.manual_seed(0)
    conf = {"data": {"image_shape": "3,32,32", "filter": "topN(2)"}, "params": {"batch_size": 8}, "log_dir": str(tmp_path)}
    model = TinyCls()
    srv = ClusterServing(conf, model=model, device="cpu").start(tensorboard=True)
    inq, outq = InputQueue(conf, broker=srv.db), OutputQueue(conf, broker=srv.db)
    rng = np.random.RandomState(0)
    imgs = {f"img{i}": rng.randint(0, 256, (40 + i, 50, 3), dtype=np.uint8) for i in range(20)}
    for uri, im in imgs.items():
        assert inq.enqueue_image(uri, im)
    assert srv.wait_until(20, 30)
    r = outq.query("img3")
    assert r.startswith("[[") and r.count("[") == 3          # [[idx,prob],[idx,prob]]
    top = eval(r)
    assert top[0][1] >= top[1][1] and 0 <= top[0][0] < 10
    assert outq.query("nope") == "{}"
    allr = outq.dequeue()
    assert set(allr) == set(imgs) and outq.dequeue() == {}
    st = srv.stats()
    assert st["records"] == 20 and st["p99_ms"] > 0
    srv.close()
    assert srv.summary.read_scalar("Total Records Number")

    # generic tensor records + raw output (empty filter) over the RESP socket
    class Scale(nn.Module):
        def forward(self, x):
            return x.float() * 2

    srv2 = ClusterServing({"data": {"filter": "", "src": "localhost:0"}, "params": {"batch_size": 4}}, model=Scale(), device="cpu",
                          serve_socket=True).start()
    q = InputQueue({}, host="127.0.0.1", port=srv2.port)
    o = OutputQueue({}, host="127.0.0.1", port=srv2.port)
    for i in range(6):
        q.enqueue_tensor(f"t{i}", np.full((3,), i, np.float32))
    assert srv2.wait_until(6, 30)
    time.sleep(0.05)
    assert o.query("t4").replace(" ", "") == "[8.,8.,8.]"
    srv2.close()


def test_client_backpressure():
    from zoo.native.rt import Broker
    from zoo.serving import InputQueue

    b = Broker(maxmemory=20000)
    q = InputQueue({}, broker=b)
    q.interval_if_error = 0
    ok = [q.enqueue_tensor(f"u{i}", np.zeros(256, np.float32)) for i in range(30)]
    assert ok[0] and not ok[-1]                 # refused once 60 % of maxmemory is used
    assert b.info()["used_memory"] <= 0.6 * 20000 + 1200
    b.close()


def test_serving_bulk_tensor_path():
    """enqueue_tensors -> packed batch read -> hset_many: results equal the per-record path."""
    from zoo.serving import ClusterServing, InputQueue, OutputQueue

    class Score(nn.Module):
        def forward(self, x):
            x = x.float()
            return torch.softmax(torch.stack([x[:, 0], x[:, 1], x[:, 0] - x[:, 1]], 1), -1)

    conf = {"data": {"filter": "topN(1)"}, "params": {"batch_size": 64}}
    srv = ClusterServing(conf, model=Score(), device="cpu").start()
    q, o = InputQueue(conf, broker=srv.db), OutputQueue(conf, broker=srv.db)
    rng = np.random.RandomState(0)
    data = rng.randint(0, 5, (300, 2)).astype(np.int64)
    assert q.enqueue_tensor("r0", data[0])
    assert q.enqueue_tensors([f"r{i}" for i in range(1, 300)], data[1:]) == 299
    assert srv.wait_until(300, 30)
    time.sleep(0.05)
    res = o.dequeue()
    assert len(res) == 300
    probs = torch.softmax(torch.tensor(np.stack([data[:, 0], data[:, 1], data[:, 0] - data[:, 1]], 1), dtype=torch.float32), -1)
    for i in (0, 1, 57, 299):
        idx, p = eval(res[f"r{i}"])[0]
        assert abs(p - float(probs[i].max())) < 1e-5 and float(probs[i][idx]) == pytest.approx(float(probs[i].max()))
    assert srv._tensor_meta is not None
    srv.close()


def test_two_engines_share_one_broker_over_the_socket():
    """The reference's deployment shape: several serving engines in one consumer group of ONE broker (ClusterServing.scala:105-140
    creates the group, each Flink / Spark task reads as a consumer).  Engine A owns the broker and listens; engine B and the
    clients talk to it through the RESP socket.  Every record is served exactly once, by either engine."""
    from zoo.native.rt import RespClient
    from zoo.serving import ClusterServing, InputQueue, OutputQueue

    class Scale(nn.Module):
        def forward(self, x):
            return x.float() * 3

    conf = {"data": {"filter": "", "src": "localhost:0"}, "params": {"batch_size": 4}}
    a = ClusterServing(conf, model=Scale(), device="cpu", serve_socket=True, consumer="engine-a").start()
    b = ClusterServing(conf, model=Scale(), device="cpu", broker=RespClient("127.0.0.1", a.port), consumer="engine-b").start()
    q, o = InputQueue({}, host="127.0.0.1", port=a.port), OutputQueue({}, host="127.0.0.1", port=a.port)
    n = 60
    for i in range(n):
        q.enqueue_tensor(f"r{i}", np.full((2,), i, np.float32))
        if i % 10 == 9:
            time.sleep(0.02)                 # let both consumers get a turn
    deadline = time.time() + 30
    while a.total_records + b.total_records < n and time.time() < deadline:
        time.sleep(0.02)
    assert a.total_records + b.total_records == n        # exactly once across the group
    assert a.total_records > 0 and b.total_records > 0   # and both engines took part
    time.sleep(0.1)
    res = o.dequeue()
    assert len(res) == n and res["r7"].replace(" ", "") == "[21.,21.]"
    b.close()
    a.close()
