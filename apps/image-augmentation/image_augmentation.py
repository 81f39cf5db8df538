"""Image augmentation gallery (reference app: apps/image-augmentation/image-augmentation.ipynb -- each ImageSet transformer applied
to one picture: Brightness, Hue, Saturation, ChannelOrder, ColorJitter, Resize, AspectScale, RandomAspectScale, ChannelNormalize,
PixelNormalize, CenterCrop, RandomCrop, FixedCrop, Filler, Expand, HFlip, and a chained random pipeline).  Results are written as
PNG files instead of being shown inline."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import cv2
import numpy as np

from zoo.feature.common import ChainedPreprocessing
from zoo.feature.image import (ImageAspectScale, ImageBrightness, ImageCenterCrop, ImageChannelNormalize, ImageChannelOrder, ImageColorJitter,
                               ImageExpand, ImageFiller, ImageFixedCrop, ImageHFlip, ImageHue, ImageRandomAspectScale, ImageRandomCrop,
                               ImageRandomPreprocessing, ImageResize, ImageSaturation, ImageSet)

out_dir = sys.argv[1] if len(sys.argv) > 1 else tempfile.mkdtemp()
np.random.seed(0)
yy, xx = np.mgrid[0:240, 0:320]
picture = np.stack([(xx * 255 / 320), (yy * 255 / 240), ((xx + yy) % 64) * 4], -1).astype(np.uint8)
src = os.path.join(out_dir, "source")
os.makedirs(src, exist_ok=True)
cv2.imwrite(os.path.join(src, "picture.png"), picture)

ops = {"brightness": ImageBrightness(0.0, 32.0), "hue": ImageHue(-18.0, 18.0), "saturation": ImageSaturation(10.0, 20.0),
       "channel_order": ImageChannelOrder(), "color_jitter": ImageColorJitter(), "resize": ImageResize(200, 200),
       "aspect_scale": ImageAspectScale(200, max_size=3000), "random_aspect_scale": ImageRandomAspectScale([100, 300], max_size=3000),
       "center_crop": ImageCenterCrop(200, 200), "random_crop": ImageRandomCrop(200, 200), "fixed_crop": ImageFixedCrop(0.0, 0.0, 200.0, 200.0, False),
       "filler": ImageFiller(0.0, 0.0, 0.5, 0.5, 255), "expand": ImageExpand(max_expand_ratio=2.0), "hflip": ImageHFlip(),
       "random_pipeline": ChainedPreprocessing([ImageRandomPreprocessing(ImageHFlip(), 0.5), ImageColorJitter(), ImageRandomCrop(180, 240)])}
for name, op in ops.items():
    out = ImageSet.read(src).transform(op)
    mat = np.asarray(out.get_image_features()[0]["mat"])
    cv2.imwrite(os.path.join(out_dir, name + ".png"), np.clip(mat, 0, 255).astype(np.uint8))
    print(f"{name:20s} -> {mat.shape}")
norm = ImageSet.read(src).transform(ImageChannelNormalize(123.0, 117.0, 104.0, 58.0, 57.0, 57.0)).get_image_features()[0]["mat"]
print("channel_normalize    -> mean %.2f std %.2f" % (np.mean(norm), np.std(norm)))
print("wrote", len(ops), "images to", out_dir)
