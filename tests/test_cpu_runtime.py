"""C++ runtime (libzoo_rt.so): broker verbs in-process and over the RESP socket, consumer groups,
blocking reads, memory guard; arena + record store; CRC32C / TFRecord framing."""
import os
import threading
import time

import numpy as np
import pytest


def test_broker_inprocess_streams_and_hashes():
    from zoo.native.rt import Broker, RespError

    b = Broker()
    assert b.ping()
    ids = [b.xadd("image_stream", {"uri": f"u{i}", "image": b"\x00\x01binary\r\n" * 3}) for i in range(5)]
    assert b.xlen("image_stream") == 5 and len(set(ids)) == 5
    b.xgroup_create("image_stream", "serving", id="0")
    with pytest.raises(RespError):
        b.xgroup_create("image_stream", "serving", id="0")
    got = b.xreadgroup("serving", "c0", {"image_stream": ">"}, count=3)
    assert got[0][0] == "image_stream" and len(got[0][1]) == 3
    assert got[0][1][0][1]["uri"] == b"u0" and got[0][1][0][1]["image"] == b"\x00\x01binary\r\n" * 3
    got2 = b.xreadgroup("serving", "c1", {"image_stream": ">"}, count=10)
    assert [e[1]["uri"] for e in got2[0][1]] == [b"u3", b"u4"]
    assert b.xreadgroup("serving", "c1", {"image_stream": ">"}, count=10) == []
    assert b.xack("image_stream", "serving", *[e[0] for e in got[0][1]]) == 3
    assert b.xtrim("image_stream", 2) == 3 and b.xlen("image_stream") == 2
    b.hset("result:u0", "value", "[[1,0.9]]")
    b.hset("result:u1", mapping={"value": "x", "extra": "y"})
    assert b.hgetall("result:u0") == {b"value": b"[[1,0.9]]"}
    assert sorted(b.keys("result:*")) == [b"result:u0", b"result:u1"]
    assert b.delete("result:u0") == 1 and b.keys("result:*") == [b"result:u1"]
    info = b.info("memory")
    assert info["used_memory"] > 0 and info["maxmemory"] == 0
    b.close()


def test_broker_memory_guard_and_blocking_read():
    from zoo.native.rt import Broker, RespError

    b = Broker(maxmemory=4096)
    with pytest.raises(RespError, match="OOM"):
        for i in range(100):
            b.xadd("s", {"image": b"x" * 512})
    b.config_set("maxmemory", "1mb")
    assert b.info()["maxmemory"] == 1 << 20
    b.flushall()
    out = {}

    def reader():
        out["r"] = b.xread({"q": "$"}, count=1, block=2000)

    t = threading.Thread(target=reader)
    t.start()
    time.sleep(0.1)
    b.xadd("q", {"k": "v"})
    t.join(3)
    assert out["r"] and out["r"][0][1][0][1] == {"k": b"v"}
    t0 = time.time()
    assert b.xread({"q": "$"}, block=50) == []
    assert 0.03 < time.time() - t0 < 1.0
    b.close()


def test_broker_resp_socket_roundtrip():
    from zoo.native.rt import Broker, RespClient

    b = Broker()
    port = b.serve("127.0.0.1", 0)
    c = RespClient("127.0.0.1", port)
    assert c.ping()
    big = os.urandom(300_000)
    c.xadd("image_stream", {"uri": "a", "image": big})
    assert b.xlen("image_stream") == 1                     # same object in-process
    b.xgroup_create("image_stream", "g", id="0")
    r = c.xreadgroup("g", "c", {"image_stream": ">"}, count=4, block=100)
    assert r[0][1][0][1]["image"] == big
    b.hset("result:a", "value", "ok")
    assert c.hgetall("result:a") == {b"value": b"ok"}
    cs = [RespClient("127.0.0.1", port) for _ in range(4)]   # several concurrent connections

    def push(cl, k):
        for i in range(50):
            cl.xadd("multi", {"i": f"{k}-{i}"})

    ts = [threading.Thread(target=push, args=(cl, k)) for k, cl in enumerate(cs)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert c.xlen("multi") == 200
    for cl in cs + [c]:
        cl.close()
    b.stop_server()
    b.close()


def test_arena_and_record_store(tmp_path):
    from zoo.native.rt import Arena, RecordStore

    a = Arena(1 << 20)
    x = a.ndarray((100, 7), np.float32)
    x[:] = np.arange(700, dtype=np.float32).reshape(100, 7)
    off = a.alloc(10)
    assert off % 64 == 0 and a.used >= 2800
    with pytest.raises(MemoryError):
        a.alloc(2 << 20)
    rs = RecordStore(a)
    recs = [os.urandom(n) for n in (1, 17, 4096, 300, 0, 65)]
    for r in recs:
        rs.append(r)
    assert len(rs) == 6 and rs.get(2) == recs[2] and rs.get(4) == b""
    buf, offs = rs.gather([5, 0, 2])
    assert bytes(buf[offs[0]:offs[1]]) == recs[5] and bytes(buf[offs[2]:offs[3]]) == recs[2] and offs[-1] == 65 + 1 + 4096
    with pytest.raises(IndexError):
        rs.get(99)
    np.testing.assert_array_equal(x[3], np.arange(21, 28, dtype=np.float32))     # untouched by later allocations
    rs.close(); a.close()
    f = Arena(1 << 16, kind="file", path=str(tmp_path / "slab.bin"))
    v = f.ndarray((16,), np.int64)
    v[:] = 7
    assert f.evict(0, 1 << 16) == 0
    assert int(v.sum()) == 7 * 16                            # pages come back from the file
    f.close()
    assert os.path.getsize(tmp_path / "slab.bin") == 1 << 16


def test_crc_and_tfrecord_match_python_writer():
    from zoo.native import rt
    from zoo.tensorboard import writer as w

    for data in (b"", b"123456789", os.urandom(1000)):
        assert rt.crc32c(data) == w.crc32c(data)
        assert rt.masked_crc32c(data) == w.masked_crc(data)
        assert rt.tfrecord_frame(data) == w.frame_record(data)
    assert rt.crc32c(b"123456789") == 0xE3069283


def test_broker_native_topn_results_match_python_format():
    import ast

    import numpy as np

    from zoo.native.rt import Broker
    from zoo.ops.serving_ops import format_topn

    b = Broker()
    idx = np.array([[3, 1], [0, 2], [7, 7]], np.int32)
    val = np.array([[0.75, 0.125], [1e-8, 0.3], [0.99999994, 123456.79]], np.float32)
    assert b.hset_topn("result:", "value", b"a\nb/c\nlast\n", idx, val) == 3
    for k, uri in enumerate(["a", "b/c", "last"]):
        got = b.hgetall("result:" + uri)
        got = {(k2.decode() if isinstance(k2, bytes) else k2): (v.decode() if isinstance(v, bytes) else v) for k2, v in got.items()}["value"]
        parsed, want = ast.literal_eval(got), ast.literal_eval(format_topn(idx[k], val[k]))
        assert [p[0] for p in parsed] == [w[0] for w in want]
        np.testing.assert_allclose([p[1] for p in parsed], [w[1] for w in want], rtol=1e-7)     # shortest round-trip float32 digits
        assert np.float32(parsed[0][1]) == val[k, 0]
    b.config_set("maxmemory", 1)
    import pytest

    from zoo.native.rt import RespError

    with pytest.raises(RespError):
        b.hset_topn("result:", "value", b"zzz\n", idx[:1], val[:1])
    b.close()


def test_broker_socket_rejects_malformed_requests():
    """A request that is not RESP (or is oversized) is answered with -ERR and the connection is closed; the broker keeps serving."""
    import socket

    from zoo.native import rt

    b = rt.Broker()
    port = b.serve("127.0.0.1", 0)
    try:
        for bad in (b"*2\r\n$3\r\nGET\r\nnot-a-bulk-header\r\n", b"*99999999999\r\n", b"*1\r\n$-5x\r\n", b"*1\r\n$4\r\nPINGxx"):
            c = socket.create_connection(("127.0.0.1", port), timeout=5)
            c.sendall(bad)
            data = c.recv(4096)
            assert data.startswith(b"-ERR"), (bad, data)
            assert c.recv(16) == b""            # closed by the server
            c.close()
        c = socket.create_connection(("127.0.0.1", port), timeout=5)
        c.sendall(b"*1\r\n$4\r\nPING\r\n")
        assert c.recv(64).startswith(b"+PONG")
        c.sendall(b"*5\r\n$4\r\nXADD\r\n$1\r\ns\r\n$1\r\n*\r\n$3\r\nuri\r\n$3\r\na\nb\r\n")
        assert c.recv(128).startswith(b"-ERR")
        c.close()
    finally:
        b.stop_server()


def test_runtime_stress_harness_under_thread_sanitizer():
    """The C++ runtime (broker incl. its RESP server, arena, record store) under ThreadSanitizer with the multi-threaded stress
    harness tests/native/rt_stress.cpp (summaries from dev/tsan_runtime.sh are kept in profiles/sanitizer_runtime_*_r2.txt)."""
    import shutil
    import subprocess

    if shutil.which("g++") is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = "/tmp/rt_stress_tsan_test"
    build = subprocess.run(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-pthread", "csrc/runtime/arena.cpp", "csrc/runtime/broker.cpp",
                            "tests/native/rt_stress.cpp", "-o", exe], cwd=root, capture_output=True, text=True)
    if build.returncode != 0 and "sanitize" in build.stderr:
        pytest.skip("this toolchain has no ThreadSanitizer runtime")
    assert build.returncode == 0, build.stderr[-2000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert run.returncode == 0 and "RT_STRESS_OK" in run.stdout and "WARNING: ThreadSanitizer" not in run.stderr, run.stdout[-1000:] + run.stderr[-3000:]
