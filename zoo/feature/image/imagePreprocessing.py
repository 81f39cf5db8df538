"""Image preprocessing ops over ImageFeatures (reference: pyzoo/zoo/feature/image/imagePreprocessing.py:25-390,
zoo/.../feature/image/*.scala -- OpenCV on the CPU).  Each op reads/writes the `mat` (HWC, BGR) of a feature; the
random ops draw from numpy's global RNG.  For whole batches on the GPU see `zoo.ops.serving_ops.preprocess_images`
(kernel K13: resize + crop + normalise + layout in one pass)."""
from __future__ import annotations

import os

import numpy as np

from ..common import ChainedPreprocessing, Preprocessing, Sample  # noqa: F401


class ImagePreprocessing(Preprocessing):
    def __init__(self, bigdl_type="float", *args):
        pass

    def apply(self, f):
        f["mat"] = self.transform_mat(np.asarray(f["mat"]), f)
        return f

    def transform_mat(self, mat, f):  # pragma: no cover - interface
        return mat


class ImageBytesToMat(ImagePreprocessing):
    def __init__(self, byte_key="bytes", image_codec=-1, bigdl_type="float"):
        self.byte_key, self.image_codec = byte_key, image_codec

    def apply(self, f):
        import cv2

        mat = cv2.imdecode(np.frombuffer(f[self.byte_key], np.uint8), cv2.IMREAD_COLOR if self.image_codec in (-1, 1) else self.image_codec)
        if mat is None:
            raise ValueError(f"cannot decode image {f.get('uri')}")
        f["mat"], f["originalSize"] = mat, mat.shape
        return f


class ImagePixelBytesToMat(ImagePreprocessing):
    """Raw pixel bytes (h*w*c uint8, shape under `originalSize`) -> mat."""

    def __init__(self, byte_key="bytes", bigdl_type="float"):
        self.byte_key = byte_key

    def apply(self, f):
        h, w, c = f["originalSize"]
        f["mat"] = np.frombuffer(f[self.byte_key], np.uint8).reshape(h, w, c)
        return f


class ImageResize(ImagePreprocessing):
    def __init__(self, resize_h, resize_w, resize_mode=1, use_scale_factor=True, bigdl_type="float"):
        self.h, self.w, self.mode = int(resize_h), int(resize_w), resize_mode

    def transform_mat(self, mat, f):
        import cv2

        interp = {0: cv2.INTER_NEAREST, 1: cv2.INTER_LINEAR, 2: cv2.INTER_CUBIC, 3: cv2.INTER_AREA, 4: cv2.INTER_LANCZOS4}.get(self.mode, cv2.INTER_LINEAR)
        if self.mode == -1:
            interp = np.random.choice([cv2.INTER_LINEAR, cv2.INTER_CUBIC, cv2.INTER_AREA, cv2.INTER_NEAREST, cv2.INTER_LANCZOS4])
        return cv2.resize(mat, (self.w, self.h), interpolation=int(interp))


class ImageBrightness(ImagePreprocessing):
    def __init__(self, delta_low, delta_high, bigdl_type="float"):
        self.lo, self.hi = float(delta_low), float(delta_high)

    def transform_mat(self, mat, f):
        return np.clip(mat.astype(np.float32) + np.random.uniform(self.lo, self.hi), 0, 255).astype(mat.dtype)


class ImageChannelNormalize(ImagePreprocessing):
    """(pixel - mean) / std per channel; means are given in R, G, B order and applied to the BGR mat."""

    def __init__(self, mean_r, mean_g, mean_b, std_r=1.0, std_g=1.0, std_b=1.0, bigdl_type="float"):
        self.mean = np.array([mean_b, mean_g, mean_r], dtype=np.float32)
        self.std = np.array([std_b, std_g, std_r], dtype=np.float32)

    def transform_mat(self, mat, f):
        return (mat.astype(np.float32) - self.mean) / self.std


class PerImageNormalize(ImagePreprocessing):
    """Min-max normalise each image into [min, max] (cv2.NORM_MINMAX = 32)."""

    def __init__(self, min, max, norm_type=32, bigdl_type="float"):
        self.lo, self.hi = float(min), float(max)

    def transform_mat(self, mat, f):
        m = mat.astype(np.float32)
        a, b = m.min(), m.max()
        return (m - a) / max(b - a, 1e-12) * (self.hi - self.lo) + self.lo


class ImageMatToTensor(ImagePreprocessing):
    """mat (HWC) -> CHW float tensor under `tensor_key` (`format="NHWC"` keeps HWC)."""

    def __init__(self, to_RGB=False, tensor_key="imageTensor", share_buffer=True, format="NCHW", bigdl_type="float"):
        self.to_rgb, self.key, self.format = to_RGB, tensor_key, format

    def apply(self, f):
        m = np.asarray(f["mat"], dtype=np.float32)
        if m.ndim == 2:
            m = m[:, :, None]
        if self.to_rgb and m.shape[2] >= 3:
            m = m[:, :, [2, 1, 0] + list(range(3, m.shape[2]))]
        f[self.key] = np.ascontiguousarray(m.transpose(2, 0, 1) if self.format == "NCHW" else m)
        return f


class ImageSetToSample(ImagePreprocessing):
    def __init__(self, input_keys=("imageTensor",), target_keys=("label",), sample_key="sample", bigdl_type="float"):
        self.input_keys, self.target_keys, self.sample_key = list(input_keys), list(target_keys or []), sample_key

    def apply(self, f):
        xs = [np.asarray(f[k], dtype=np.float32) for k in self.input_keys]
        ys = [np.asarray(f[k], dtype=np.float32) for k in self.target_keys if k in f]
        f[self.sample_key] = Sample(xs[0] if len(xs) == 1 else xs, (ys[0] if len(ys) == 1 else ys) if ys else None)
        return f


def _hsv_op(mat, fn):
    import cv2

    hsv = cv2.cvtColor(np.clip(mat, 0, 255).astype(np.uint8), cv2.COLOR_BGR2HSV).astype(np.float32)
    hsv = fn(hsv)
    hsv[..., 0] = np.mod(hsv[..., 0], 180)
    hsv[..., 1:] = np.clip(hsv[..., 1:], 0, 255)
    return cv2.cvtColor(hsv.astype(np.uint8), cv2.COLOR_HSV2BGR)


class ImageHue(ImagePreprocessing):
    def __init__(self, delta_low, delta_high, bigdl_type="float"):
        self.lo, self.hi = float(delta_low), float(delta_high)

    def transform_mat(self, mat, f):
        d = np.random.uniform(self.lo, self.hi)

        def fn(h):
            h[..., 0] += d / 2.0          # OpenCV hue is in [0, 180)
            return h

        return _hsv_op(mat, fn)


class ImageSaturation(ImagePreprocessing):
    def __init__(self, delta_low, delta_high, bigdl_type="float"):
        self.lo, self.hi = float(delta_low), float(delta_high)

    def transform_mat(self, mat, f):
        s = np.random.uniform(self.lo, self.hi)

        def fn(h):
            h[..., 1] *= s
            return h

        return _hsv_op(mat, fn)


class ImageChannelOrder(ImagePreprocessing):
    """Random permutation of the colour channels."""

    def transform_mat(self, mat, f):
        return mat[:, :, np.random.permutation(mat.shape[2])]


class ImageContrast(ImagePreprocessing):
    def __init__(self, delta_low, delta_high, bigdl_type="float"):
        self.lo, self.hi = float(delta_low), float(delta_high)

    def transform_mat(self, mat, f):
        return np.clip(mat.astype(np.float32) * np.random.uniform(self.lo, self.hi), 0, 255).astype(mat.dtype)


class ImageColorJitter(ImagePreprocessing):
    """Random brightness / contrast / saturation / hue (each with its own probability), optionally in random order,
    then an optional random channel permutation (ColorJitter.scala)."""

    def __init__(self, brightness_prob=0.5, brightness_delta=32.0, contrast_prob=0.5, contrast_lower=0.5, contrast_upper=1.5, hue_prob=0.5,
                 hue_delta=18.0, saturation_prob=0.5, saturation_lower=0.5, saturation_upper=1.5, random_order_prob=0.0, shuffle=False,
                 bigdl_type="float"):
        self.ops = [(brightness_prob, ImageBrightness(-brightness_delta, brightness_delta)), (contrast_prob, ImageContrast(contrast_lower, contrast_upper)),
                    (saturation_prob, ImageSaturation(saturation_lower, saturation_upper)), (hue_prob, ImageHue(-hue_delta, hue_delta))]
        self.random_order_prob, self.shuffle = random_order_prob, shuffle

    def transform_mat(self, mat, f):
        ops = list(self.ops)
        if self.shuffle or np.random.rand() < self.random_order_prob:
            np.random.shuffle(ops)
        for p, op in ops:
            if np.random.rand() < p:
                mat = op.transform_mat(mat, f)
        return mat


class ImageAspectScale(ImagePreprocessing):
    """Resize so the short side is `min_size` (long side capped by `max_size`), keeping the aspect ratio."""

    def __init__(self, min_size, scale_multiple_of=1, max_size=1000, resize_mode=1, use_scale_factor=True, min_scale=-1.0, bigdl_type="float"):
        self.min_size, self.mult, self.max_size, self.min_scale = int(min_size), int(scale_multiple_of), int(max_size), float(min_scale)

    def _target(self, h, w, min_size):
        s = min_size / min(h, w)
        if self.min_scale > 0:
            s = max(s, self.min_scale)
        if round(s * max(h, w)) > self.max_size:
            s = self.max_size / max(h, w)
        nh, nw = int(round(h * s)), int(round(w * s))
        if self.mult > 1:
            nh, nw = max(self.mult, nh // self.mult * self.mult), max(self.mult, nw // self.mult * self.mult)
        return nh, nw

    def transform_mat(self, mat, f):
        import cv2

        nh, nw = self._target(mat.shape[0], mat.shape[1], self.min_size)
        return cv2.resize(mat, (nw, nh))


class ImageRandomAspectScale(ImageAspectScale):
    def __init__(self, scales, scale_multiple_of=1, max_size=1000, bigdl_type="float"):
        super().__init__(scales[0], scale_multiple_of, max_size)
        self.scales = list(scales)

    def transform_mat(self, mat, f):
        import cv2

        nh, nw = self._target(mat.shape[0], mat.shape[1], int(np.random.choice(self.scales)))
        return cv2.resize(mat, (nw, nh))


class ImagePixelNormalize(ImagePreprocessing):
    """Subtract a per-pixel mean image (flattened HWC `means`)."""

    def __init__(self, means, bigdl_type="float"):
        self.means = np.asarray(means, dtype=np.float32)

    def transform_mat(self, mat, f):
        return mat.astype(np.float32) - self.means.reshape(mat.shape)


class _Crop(ImagePreprocessing):
    def _crop(self, mat, x1, y1, x2, y2, is_clip=True):
        h, w = mat.shape[:2]
        if is_clip:
            x1, y1, x2, y2 = max(0, x1), max(0, y1), min(w, x2), min(h, y2)
        return mat[int(y1):int(y2), int(x1):int(x2)]


class ImageRandomCrop(_Crop):
    def __init__(self, crop_width, crop_height, is_clip=True, bigdl_type="float"):
        self.cw, self.ch, self.is_clip = int(crop_width), int(crop_height), is_clip

    def transform_mat(self, mat, f):
        h, w = mat.shape[:2]
        x = np.random.randint(0, max(w - self.cw, 0) + 1)
        y = np.random.randint(0, max(h - self.ch, 0) + 1)
        return self._crop(mat, x, y, x + self.cw, y + self.ch, self.is_clip)


class ImageCenterCrop(_Crop):
    def __init__(self, crop_width, crop_height, is_clip=True, bigdl_type="float"):
        self.cw, self.ch, self.is_clip = int(crop_width), int(crop_height), is_clip

    def transform_mat(self, mat, f):
        h, w = mat.shape[:2]
        x, y = (w - self.cw) // 2, (h - self.ch) // 2
        return self._crop(mat, x, y, x + self.cw, y + self.ch, self.is_clip)


class ImageFixedCrop(_Crop):
    def __init__(self, x1, y1, x2, y2, normalized=True, is_clip=True, bigdl_type="float"):
        self.box, self.normalized, self.is_clip = (float(x1), float(y1), float(x2), float(y2)), normalized, is_clip

    def transform_mat(self, mat, f):
        h, w = mat.shape[:2]
        x1, y1, x2, y2 = self.box
        if self.normalized:
            x1, x2, y1, y2 = x1 * w, x2 * w, y1 * h, y2 * h
        return self._crop(mat, round(x1), round(y1), round(x2), round(y2), self.is_clip)


class ImageExpand(ImagePreprocessing):
    """Place the image at a random position inside a larger canvas filled with the channel means (SSD augmentation)."""

    def __init__(self, means_r=123, means_g=117, means_b=104, min_expand_ratio=1.0, max_expand_ratio=4.0, bigdl_type="float"):
        self.means, self.lo, self.hi = (means_b, means_g, means_r), float(min_expand_ratio), float(max_expand_ratio)

    def transform_mat(self, mat, f):
        r = np.random.uniform(self.lo, self.hi)
        h, w = mat.shape[:2]
        nh, nw = int(h * r), int(w * r)
        y, x = np.random.randint(0, nh - h + 1), np.random.randint(0, nw - w + 1)
        out = np.empty((nh, nw, mat.shape[2]), dtype=mat.dtype)
        out[:] = np.asarray(self.means[:mat.shape[2]], dtype=mat.dtype)
        out[y:y + h, x:x + w] = mat
        f["expandBox"] = (x, y, nw, nh)
        return out


class ImageFiller(ImagePreprocessing):
    """Fill the (normalised) rectangle with `value`."""

    def __init__(self, start_x, start_y, end_x, end_y, value=255, bigdl_type="float"):
        self.box, self.value = (float(start_x), float(start_y), float(end_x), float(end_y)), value

    def transform_mat(self, mat, f):
        h, w = mat.shape[:2]
        x1, y1, x2, y2 = self.box
        out = mat.copy()
        out[int(np.ceil(y1 * h)):int(np.ceil(y2 * h)), int(np.ceil(x1 * w)):int(np.ceil(x2 * w))] = self.value
        return out


class ImageHFlip(ImagePreprocessing):
    def transform_mat(self, mat, f):
        return mat[:, ::-1].copy()


ImageMirror = ImageHFlip


class ImageFeatureToTensor(Preprocessing):
    def __init__(self, bigdl_type="float"):
        pass

    def apply(self, f):
        return f.get_tensor()


class ImageFeatureToSample(Preprocessing):
    def __init__(self, bigdl_type="float"):
        pass

    def apply(self, f):
        return f["sample"] if "sample" in f else Sample(f.get_tensor(), f.get("label"))


class RowToImageFeature(Preprocessing):
    """A DataFrame image row (`NNImageReader` struct: origin, height, width, nChannels, mode, data) -> ImageFeature."""

    def __init__(self, bigdl_type="float"):
        pass

    def apply(self, row):
        from .imageset import ImageFeature

        r = row["image"] if isinstance(row, dict) and "image" in row else row
        data = np.asarray(r["data"])
        mat = data.reshape(r["height"], r["width"], r["nChannels"]) if data.ndim == 1 else data
        return ImageFeature(uri=r.get("origin"), mat=mat, originalSize=mat.shape)


class ImageRandomPreprocessing(Preprocessing):
    """Apply `preprocessing` with probability `prob`."""

    def __init__(self, preprocessing, prob, bigdl_type="float"):
        self.preprocessing, self.prob = preprocessing, float(prob)

    def apply(self, f):
        return self.preprocessing(f) if np.random.rand() < self.prob else f


ImagePixelNormalizer = ImagePixelNormalize      # the Scala name (feature/image/ImagePixelNormalizer.scala)
ImageProcessing = ImagePreprocessing            # the Scala base-class name (feature/image/ImageProcessing.scala)


class ImageChannelScaledNormalizer(ImagePreprocessing):
    """`(pixel - channel mean) * scale` in one pass (reference: feature/image/ImageChannelScaledNormalizer.scala:25-50; the means
    are given as R, G, B while the mat is BGR)."""

    def __init__(self, mean_r, mean_g, mean_b, scale, bigdl_type="float"):
        self.mean = np.array([mean_b, mean_g, mean_r], dtype=np.float32)
        self.scale = np.float32(scale)

    def transform_mat(self, mat, f):
        return (mat.astype(np.float32) - self.mean[:mat.shape[2]]) * self.scale


class ImageRandomCropper(_Crop):
    """Crop `crop_width x crop_height` at a random (`cropper_method="random"`) or the centre (`"center"`) position and, when
    `mirror` is set, flip the crop horizontally with probability 1/2 (feature/image/ImageRandomCropper.scala:27-62)."""

    def __init__(self, crop_width, crop_height, mirror, cropper_method="random", channels=3, bigdl_type="float"):
        method = str(cropper_method).lower()
        if method not in ("random", "center", "croprandom", "cropcenter"):
            raise ValueError(f"cropper_method {cropper_method!r}: expected 'random' or 'center'")
        self.cw, self.ch, self.mirror, self.center, self.channels = int(crop_width), int(crop_height), bool(mirror), "center" in method, channels

    def transform_mat(self, mat, f):
        h, w = mat.shape[:2]
        if h < self.ch or w < self.cw:
            raise ValueError(f"image {h}x{w} is smaller than the crop {self.ch}x{self.cw}")
        if self.center:
            x, y = (w - self.cw) // 2, (h - self.ch) // 2
        else:
            x, y = np.random.randint(0, w - self.cw + 1), np.random.randint(0, h - self.ch + 1)
        out = self._crop(mat, x, y, x + self.cw, y + self.ch)
        return out[:, ::-1].copy() if self.mirror and np.random.rand() < 0.5 else out


class ImageRandomResize(ImagePreprocessing):
    """Resize so that the SHORTER side becomes a size drawn uniformly from [min_size, max_size], keeping the aspect ratio
    (feature/image/ImageRandomResize.scala; the Inception / ResNet training pipelines use it before the random crop)."""

    def __init__(self, min_size, max_size, bigdl_type="float"):
        if min_size > max_size:
            raise ValueError("min_size must not exceed max_size")
        self.min_size, self.max_size = int(min_size), int(max_size)

    def transform_mat(self, mat, f):
        import cv2

        h, w = mat.shape[:2]
        short = np.random.randint(self.min_size, self.max_size + 1)
        nh, nw = (short, max(1, int(round(w * short / h)))) if h <= w else (max(1, int(round(h * short / w))), short)
        return cv2.resize(mat, (nw, nh), interpolation=cv2.INTER_LINEAR)


class ImageMatToFloats(ImagePreprocessing):
    """mat -> flat float array under `out_key` (HWC order), with the size kept under "originalSize"-style keys
    (feature/image/ImageMatToFloats.scala:28-50).  `valid_*` size the reusable buffer when `share_buffer` is on: with a shared
    buffer the array is only valid until the next feature passes through, as in the reference."""

    def __init__(self, valid_height=300, valid_width=300, valid_channel=3, out_key="floats", share_buffer=True, bigdl_type="float"):
        self.out_key, self.share = out_key, bool(share_buffer)
        self._buf = np.empty(int(valid_height) * int(valid_width) * int(valid_channel), dtype=np.float32) if self.share else None

    def apply(self, f):
        mat = np.asarray(f["mat"])
        n = mat.size
        if self.share:
            if self._buf.size < n:
                self._buf = np.empty(n, dtype=np.float32)
            out = self._buf[:n]
            np.copyto(out, mat.reshape(-1), casting="unsafe")
        else:
            out = mat.astype(np.float32).reshape(-1)
        f[self.out_key] = out
        f["size"] = mat.shape
        return f


class BufferedImageResize(ImagePreprocessing):
    """Resize the ENCODED image under `bytes` and re-encode it in its own format, taken from the uri's extension
    (feature/image/BufferedImageResize.scala:27-50; there through java.awt with SCALE_SMOOTH, here OpenCV INTER_AREA)."""

    def __init__(self, resize_h, resize_w, bigdl_type="float"):
        self.h, self.w = int(resize_h), int(resize_w)

    def apply(self, f):
        import cv2

        img = cv2.imdecode(np.frombuffer(f["bytes"], np.uint8), cv2.IMREAD_UNCHANGED)
        if img is None:
            raise ValueError(f"cannot decode image {f.get('uri')}")
        if img.shape[:2] != (self.h, self.w):
            ext = os.path.splitext(str(f.get("uri", "")))[1].lower() or ".png"
            ok, enc = cv2.imencode(ext if ext in (".jpg", ".jpeg", ".png", ".bmp", ".webp", ".tif", ".tiff") else ".png",
                                   cv2.resize(img, (self.w, self.h), interpolation=cv2.INTER_AREA))
            if not ok:
                raise ValueError(f"cannot encode image {f.get('uri')}")
            f["bytes"] = enc.tobytes()
        return f
