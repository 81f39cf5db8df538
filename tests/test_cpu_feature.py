"""Feature-layer additions of round 2 (image ops that wrap BigDL transforms in the reference, MultiTensorsToSample)."""
import numpy as np
import pytest


def test_image_ops_added_in_round_2():
    """ImageChannelScaledNormalizer, ImageRandomCropper, ImageRandomResize, ImageMatToFloats, BufferedImageResize,
    MultiTensorsToSample (reference: zoo/.../feature/image/*.scala of the same names, feature/common/MultiTensorsToSample.scala)."""
    import cv2

    from zoo.feature.common import MultiTensorsToSample, SeqToMultipleTensors
    from zoo.feature.image import (BufferedImageResize, ImageChannelScaledNormalizer, ImageMatToFloats, ImagePixelNormalizer,
                                   ImageRandomCropper, ImageRandomResize)

    np.random.seed(0)
    mat = np.random.randint(0, 255, (40, 60, 3)).astype(np.uint8)
    out = ImageChannelScaledNormalizer(123, 117, 104, 0.0078125)({"mat": mat.copy()})["mat"]
    assert out.dtype == np.float32 and np.allclose(out[..., 0], (mat[..., 0].astype(np.float32) - 104) * 0.0078125)    # channel 0 is B
    assert np.allclose(out[..., 2], (mat[..., 2].astype(np.float32) - 123) * 0.0078125)

    c = ImageRandomCropper(32, 24, mirror=False, cropper_method="center")({"mat": mat.copy()})["mat"]
    assert c.shape == (24, 32, 3) and np.array_equal(c, mat[8:32, 14:46])
    seen_flip = seen_plain = False
    for _ in range(20):
        r = ImageRandomCropper(60, 40, mirror=True)({"mat": mat.copy()})["mat"]          # the only possible crop: the whole image
        seen_flip |= np.array_equal(r, mat[:, ::-1])
        seen_plain |= np.array_equal(r, mat)
    assert seen_flip and seen_plain
    with pytest.raises(ValueError):
        ImageRandomCropper(100, 100, False)({"mat": mat.copy()})

    for _ in range(5):
        r = ImageRandomResize(20, 30)({"mat": mat.copy()})["mat"]
        assert 20 <= r.shape[0] <= 30 and abs(r.shape[1] / r.shape[0] - 1.5) < 0.05      # shorter side drawn, aspect kept

    op = ImageMatToFloats(40, 60, 3)
    f = op({"mat": mat.copy()})
    assert f["floats"].dtype == np.float32 and f["floats"].shape == (40 * 60 * 3,) and np.array_equal(f["floats"], mat.reshape(-1))
    g = op({"mat": mat[:10].copy()})
    assert g["floats"].base is op._buf or g["floats"].base is op._buf.base            # shared buffer
    own = ImageMatToFloats(share_buffer=False, out_key="data")({"mat": mat.copy()})
    assert own["data"].shape == (7200,)

    ok, enc = cv2.imencode(".png", mat)
    f = BufferedImageResize(20, 30)({"bytes": enc.tobytes(), "uri": "file:/x/picture.png"})
    back = cv2.imdecode(np.frombuffer(f["bytes"], np.uint8), cv2.IMREAD_COLOR)
    assert back.shape == (20, 30, 3) and f["bytes"][:4] == b"\x89PNG"
    same = BufferedImageResize(40, 60)({"bytes": enc.tobytes(), "uri": "a.png"})
    assert same["bytes"] == enc.tobytes()                                                # already that size: bytes untouched
    assert ImagePixelNormalizer is not None

    s = MultiTensorsToSample()(SeqToMultipleTensors([[2], [3]])([1.0, 2.0, 3.0, 4.0, 5.0]))
    assert [f.shape for f in s.features] == [(2,), (3,)] and s.labels == []


def test_hadoop_sequence_files_of_images(tmp_path):
    """SequenceFile container (uncompressed / record- / block-compressed) and the ImageNet record layout; the reference's
    mini_imagenet_seq fixture written by Hadoop (PersistentMemorySpec.scala:116-131 reads the same folder: 5 images -> 3 batches of 2)."""
    import os
    import struct
    import zlib

    from zoo.feature.image import ImageSet
    from zoo.feature.image import seqfile as sf

    for v in (0, 5, -5, 127, -112, 128, -113, 300, -300, 70000, -70000, 2 ** 31 - 1, -2 ** 31, 2 ** 40):
        enc = sf._write_vint(v)
        assert sf._read_vint(enc, 0) == (v, len(enc)), v
    rng = np.random.RandomState(0)
    mats = [rng.randint(0, 255, (5 + i, 7, 3)).astype(np.uint8) for i in range(4)]
    recs = [(("img%d\n%d" % (i, i + 1)).encode() if i % 2 else str(i + 1).encode(), sf.encode_image_record(m)) for i, m in enumerate(mats)]
    p = str(tmp_path / "part-00000")
    sf.write_sequence_file(p, recs, sync_every=100)                              # small sync interval: markers between records
    assert list(sf.read_sequence_file(p)) == recs
    (tmp_path / "_SUCCESS").write_bytes(b"")
    s = ImageSet.read_sequence_files(str(tmp_path), class_num=3)                 # label 4 is dropped
    feats = s.get_image_features()
    assert len(feats) == 3 and [float(f["label"][0]) for f in feats] == [1.0, 2.0, 3.0] and feats[1]["uri"] == "img1"
    assert all(np.array_equal(f["mat"], m) for f, m in zip(feats, mats))

    # record- and block-compressed forms, built by hand with the default codec
    T = sf._TEXT
    text = lambda b: sf._write_vint(len(b)) + b
    head = lambda c, blk: (b"SEQ\x06" + text(T.encode()) + text(T.encode()) + bytes([c, blk]) + (text(sf._DEFAULT_CODEC.encode()) if c else b"") +
                           struct.pack(">i", 0) + b"S" * 16)
    body = b""
    for k, v in recs:
        kk, vv = text(k), zlib.compress(text(v))
        body += struct.pack(">ii", len(kk) + len(vv), len(kk)) + kk + vv
    (tmp_path / "rc.seq").write_bytes(head(1, 0) + body)
    assert list(sf.read_sequence_file(str(tmp_path / "rc.seq"))) == recs
    ks, vs = [text(k) for k, _ in recs], [text(v) for _, v in recs]
    blocks = [b"".join(sf._write_vint(len(k)) for k in ks), b"".join(ks), b"".join(sf._write_vint(len(v)) for v in vs), b"".join(vs)]
    blk = struct.pack(">i", -1) + b"S" * 16 + sf._write_vint(len(recs)) + b"".join(sf._write_vint(len(zlib.compress(b))) + zlib.compress(b) for b in blocks)
    (tmp_path / "bc.seq").write_bytes(head(1, 1) + blk)
    assert list(sf.read_sequence_file(str(tmp_path / "bc.seq"))) == recs

    ref = "/root/reference/zoo/src/test/resources/pmem/mini_imagenet_seq"
    if os.path.exists(ref):
        s = ImageSet.read_sequence_files(ref)
        feats = s.get_image_features()
        assert len(feats) == 5 and sorted(float(f["label"][0]) for f in feats) == [91.0, 328.0, 620.0, 697.0, 733.0]
        # the generator scaled the shorter side to 256
        assert all(f["mat"].ndim == 3 and f["mat"].shape[2] == 3 and min(f["mat"].shape[:2]) == 256 for f in feats)
