"""Auxiliary subsystems: tracing helpers and the rank watchdog."""
import logging
import os
import subprocess
import sys
import time

import pytest
import torch

from zoo.utils.profiler import PhaseTimer, nvtx_range, throughputing, time_it
from zoo.utils.watchdog import HeartbeatBoard, RankWatchdog


def test_time_it_and_throughputing_log_lines(caplog):
    with caplog.at_level(logging.INFO, logger="zoo.profiler"):
        with time_it("load model"):
            time.sleep(0.01)
        with throughputing("batch", 100, loss=lambda: 0.5):
            time.sleep(0.01)
        with nvtx_range("noop"):
            pass
    text = caplog.text
    assert "load model time elapsed [0 s," in text
    assert "records/second. Loss is 0.500000." in text


def test_phase_timer_cpu():
    pt = PhaseTimer("cpu")
    for _ in range(3):
        with pt.phase("a"):
            time.sleep(0.004)
        with pt.phase("b"):
            time.sleep(0.001)
        pt.step()
    s = pt.summary()
    assert list(s) == ["a", "b"] and s["a"]["count"] == 3 and s["a"]["mean_ms"] > s["b"]["mean_ms"] > 0.5
    assert pt.steps == 3
    pt.reset()
    assert pt.summary() == {}


def test_heartbeat_board_roundtrip(tmp_path):
    p = str(tmp_path / "hb")
    b0, b1 = HeartbeatBoard(2, 0, p), HeartbeatBoard(2, 1, p)
    b1.beat(7)
    assert b0.read(1)["step"] == 7 and b0.read(1)["pid"] == os.getpid() and b0.read(0)["step"] == 0
    b1.close()
    assert b0.read(1)["state"] == 1
    b0.close(unlink=True)
    assert not os.path.exists(p)


def test_watchdog_detects_stalled_peer_and_own_hang(tmp_path):
    p = str(tmp_path / "hb")
    seen = []
    w0 = RankWatchdog(2, 0, timeout=0.3, interval=0.05, on_failure=seen.append, path=p).start()
    peer = HeartbeatBoard(2, 1, p)
    peer.beat(1)
    for i in range(8):              # rank 0 keeps stepping, rank 1 went silent after its first beat
        w0.beat(i)
        time.sleep(0.08)
    assert seen and seen[0]["kind"] == "stalled" and seen[0]["rank"] == 1
    w0.stop(unlink=False)
    # own hang: a rank that stops beating notices it itself
    seen2 = []
    w = RankWatchdog(1, 0, timeout=0.2, interval=0.05, on_failure=seen2.append, path=str(tmp_path / "hb2")).start()
    w.beat(1)
    time.sleep(0.6)
    assert seen2 and seen2[0]["kind"] == "hang"
    w.stop()


def test_watchdog_detects_dead_process(tmp_path):
    p = str(tmp_path / "hb")
    w0 = RankWatchdog(2, 0, timeout=60, interval=0.05, on_failure=lambda f: None, path=p)
    code = ("import sys; sys.path.insert(0, %r); from zoo.utils.watchdog import HeartbeatBoard; b = HeartbeatBoard(2, 1, %r, create=False); b.beat(3)"
            % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), p))
    subprocess.run([sys.executable, "-c", code], check=True, timeout=120)     # the peer beats once and exits without closing
    w0.beat(1)
    f = w0.check()
    assert f is not None and f["kind"] == "dead" and f["rank"] == 1 and f["step"] == 3
    w0.stop()


def test_engine_beats_the_watchdog(tmp_path, monkeypatch):
    import numpy as np
    import torch.nn as nn

    from zoo.common.triggers import MaxIteration
    from zoo.feature.common import FeatureSet
    from zoo.pipeline.estimator import Estimator

    monkeypatch.setenv("ZOO_WATCHDOG_TIMEOUT", "120")
    est = Estimator(nn.Linear(4, 2), optim_methods=torch.optim.SGD)
    x, y = np.random.randn(64, 4).astype(np.float32), np.random.randint(0, 2, 64)
    est.train(FeatureSet.numpy(x, y), nn.CrossEntropyLoss(), end_trigger=MaxIteration(5), batch_size=16)
    wd = est.engine.watchdog
    assert wd is not None and wd.board.read(0)["step"] >= 5 and wd.failed is None
    est.close()
    assert est.engine.watchdog is None


def test_native_arrays_over_the_arena(tmp_path):
    import numpy as np
    import pytest

    from zoo.feature.pmem import FixLenBytesArray, FloatArray, MemoryType, VarLenBytesArray, VarLenFloatsArray

    assert MemoryType.DISK_AND_DRAM(4) == "DISK_AND_DRAM(4)"
    v = VarLenBytesArray(3, 64, MemoryType.DIRECT)
    for i, b in enumerate([b"abc", b"", b"0123456789"]):
        v.set(i, b)
    assert len(v) == 3 and v.get(0) == b"abc" and v.get(1) == b"" and v.get(2) == b"0123456789"
    with pytest.raises(ValueError):
        v.set(7, b"x")
    v.free()
    with pytest.raises(RuntimeError):
        v.get(0)
    f = FloatArray(100, 4, MemoryType.PMEM, path=str(tmp_path / "pmem.bin")).pin()       # file-backed: the "pmem mount"
    for i in range(100):
        f.set(i, [i, i + 0.5, -i, 1.0])
    np.testing.assert_allclose(f.get(7), [7, 7.5, -7, 1.0])
    assert f.batch([1, 3]).shape == (2, 4) and float(f.batch([1, 3])[1, 1]) == 3.5
    f.free()
    fx = FixLenBytesArray(4, 8)
    fx.set(2, b"12345678")
    assert fx.get(2) == b"12345678"
    with pytest.raises(ValueError):
        fx.set(0, b"short")
    vf = VarLenFloatsArray(2, 10)
    vf.set(0, [1.0, 2.0, 3.0]); vf.set(1, [4.0])
    np.testing.assert_allclose(vf.get(0), [1, 2, 3]); np.testing.assert_allclose(vf.get(1), [4])
    with pytest.raises(ValueError):
        FloatArray(1, memory_type="TAPE")


def test_spark_runner_pack_penv(tmp_path):
    """SparkRunner.pack_penv (reference: pyzoo/zoo/util/spark.py:100-113 conda-packs the environment for the executors): the archive a
    worker image needs -- the package with its native libraries, launch scripts, a lock file of the installed distributions."""
    import tarfile

    from zoo.util.spark import SparkRunner

    out = SparkRunner().pack_penv(output=str(tmp_path / "penv.tar.gz"))
    with tarfile.open(out) as t:
        names = set(t.getnames())
        lock = t.extractfile("requirements.lock").read().decode()
    assert "zoo/common/nncontext.py" in names and "scripts/zoo-submit" in names and "requirements.lock" in names
    assert not any("__pycache__" in n for n in names)
    assert any(line.startswith("torch==") for line in lock.splitlines())


def test_zoo_dictionary_and_tensor_operation(tmp_path):
    """ZooDictionary (common/ZooDictionary.scala:27-140) and TensorOperation (common/TensorOperation.scala:25-110)."""
    import torch

    from zoo.common.dictionary import ZooDictionary
    from zoo.common.tensor_operation import TensorOperation, div_tensor, expand_size, sub_tensor

    sentences = [["the", "cat", "sat"], ["the", "dog", "sat"], ["a", "cat", "ran", "the", "end"]]
    d = ZooDictionary(sentences, vocab_size=3, seed=0)
    assert d.vocabulary() == ["the", "cat", "sat"] and d.get_vocab_size() == 3 and d.get_discard_size() == 4
    assert d.get_index("the") == 0 and d.get_index("zebra") == 3                 # unknown word -> vocab_size
    assert d.get_word(1) == "cat" and d.get_word(99) in d.discard_vocab()        # unknown index -> a discarded word
    assert d.add_word("dog") == 3 and d.getIndex("dog") == 3 and "dog" not in d.discardVocab()
    assert d.encode(["the", "dog", "flew"]) == [0, 3, 4] and d.decode([0, 3]) == ["the", "dog"]
    d.save(str(tmp_path / "dict.json"))
    r = ZooDictionary.load(str(tmp_path / "dict.json"))
    assert r.word2index() == d.word2index() and r.discard_vocab() == d.discard_vocab()
    m = ZooDictionary(index2word={0: "x", 1: "y"})
    assert m.get_index("y") == 1 and m.get_word(7) in ("x", "y")

    assert expand_size(torch.zeros(2, 3, 4), torch.zeros(3, 1)) == [2, 3, 4]
    with pytest.raises(ValueError):
        expand_size(torch.zeros(2, 3), torch.zeros(4))
    x, mean = torch.arange(24.0).view(2, 3, 4), torch.arange(24.0).view(2, 3, 4).mean(-1, keepdim=True)
    want = x.clone() - mean
    assert sub_tensor(x, mean) is x and torch.equal(x, want)
    std = want.pow(2).mean(-1, keepdim=True).sqrt()
    assert torch.allclose(div_tensor(x, std), want / std)
    small = torch.ones(3, 1)
    TensorOperation.subTensor(small, torch.arange(4.0))                          # the first argument grows to the broadcast size
    assert small.shape == (3, 4) and torch.equal(small[1], 1 - torch.arange(4.0))


def test_hdf5_reader_writer_keras_weights_and_matlab_fixture(tmp_path):
    """zoo.util.hdf5: classic-layout writer / reader round trip (groups, attributes with string arrays, contiguous datasets), Keras-layout
    weight files through KerasModel.save_weights / load_weights and Net.load_keras, and the reference's MATLAB v7.3 volume
    (image3d/a.mat: chunked + deflate) through the 3-D image ops (pyzoo/test/zoo/feature/image3d/test_image3d.py)."""
    import json

    import numpy as np
    import torch

    from zoo.util.hdf5 import File, write_hdf5

    rng = np.random.RandomState(0)
    tree = {"g": {"__attrs__": {"names": np.array([b"a", b"bb"]), "scale": 2.5}, "a": rng.rand(3, 4).astype(np.float32), "sub": {"bb": np.arange(6,
            dtype=np.int64).reshape(2, 3)}},
            "top": rng.rand(5)}
    p = write_hdf5(str(tmp_path / "t.h5"), tree, attrs={"title": "zoo", "n": np.int32(3)})
    with File(p) as f:
        assert sorted(f.keys()) == ["g", "top"] and f.attrs["title"] == b"zoo" and int(f.attrs["n"]) == 3
        assert list(f["g"].attrs["names"]) == [b"a", b"bb"] and float(f["g"].attrs["scale"]) == 2.5
        assert np.array_equal(f["g/a"][...], tree["g"]["a"]) and f["g/sub/bb"][1, 2] == 5 and f["top"].shape == (5,) and f["top"].dtype == np.float64
        assert sorted(f.visit_datasets()) == ["g/a", "g/sub/bb", "top"] and "missing" not in f

    from zoo.pipeline.api.keras.layers import Dense
    from zoo.pipeline.api.keras.models import Sequential
    from zoo.pipeline.api.net import Net
    from zoo.tfpark import KerasModel

    def build():
        m = Sequential()
        m.add(Dense(5, activation="relu", input_shape=(4,)))
        m.add(Dense(2))
        m.compile(optimizer="sgd", loss="mse")
        return m

    torch.manual_seed(0)
    a, b = KerasModel(build()), KerasModel(build())
    a.save_weights(str(tmp_path / "w.h5"))
    with File(str(tmp_path / "w.h5")) as f:
        assert len(f.attrs["layer_names"]) == 2
    b.load_weights(str(tmp_path / "w.h5"))
    x = torch.randn(3, 4)
    assert torch.allclose(a.model(x), b.model(x))
    # a file in Keras' own naming (kernels [in, out]) + the json definition -> Net.load_keras
    ws = [p.detach().numpy() for p in a.model.parameters()]
    write_hdf5(str(tmp_path / "k.h5"), {"dense_1": {"__attrs__": {"weight_names": np.array([b"dense_1_W", b"dense_1_b"])},
            "dense_1_W": ws[0].T.copy(), "dense_1_b": ws[1]},
                                        "dense_2": {"__attrs__": {"weight_names": np.array([b"dense_2_W", b"dense_2_b"])},
                                                "dense_2_W": ws[2].T.copy(), "dense_2_b": ws[3]}},
               attrs={"layer_names": np.array([b"dense_1", b"dense_2"])})
    spec = {"class_name": "Sequential", "config": [{"class_name": "Dense", "config": {"output_dim": 5, "activation": "relu", "batch_input_shape": [None, 4]}},
                                                   {"class_name": "Dense", "config": {"output_dim": 2}}]}
    (tmp_path / "k.json").write_text(json.dumps(spec))
    loaded = Net.load_keras(str(tmp_path / "k.json"), str(tmp_path / "k.h5"))
    assert torch.allclose(loaded(x), a.model(x), atol=1e-6)

    mat = "/root/reference/pyzoo/test/zoo/resources/image3d/a.mat"
    if os.path.exists(mat):
        from zoo.feature.image3d.transformation import CenterCrop3D, Crop3D, Rotate3D

        with File(mat) as f:
            ds = f["meniscus_im"]
            assert ds.shape == (20, 256, 256) and ds.attrs["MATLAB_class"] == b"double"
            vol = np.asarray(ds)
        assert 0 < (vol > 0).sum() < vol.size // 100 and float(vol.max()) < 1.0           # a small segmented structure
        sample = vol[..., None]
        crop = Crop3D(start=[13, 80, 125], patch_size=[5, 40, 40])({"mat": sample, "imageTensor": sample})
        out = crop.get("imageTensor", crop.get("mat"))
        assert np.asarray(out).shape == (5, 40, 40, 1) and np.allclose(np.asarray(out)[..., 0], vol[13:18, 80:120, 125:165].astype(np.float32))
