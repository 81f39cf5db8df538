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
