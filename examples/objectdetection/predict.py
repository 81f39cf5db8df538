"""Object detection over a folder of images: ImageSet.read -> ObjectDetector.predict_image_set -> Visualizer -> image files
(reference examples: pyzoo/zoo/examples/objectdetection/predict.py and zoo/.../examples/objectdetection/inference/Predict.scala:
`ObjectDetector.load_model(path)`, `model.predict_image_set(image_set)`, `Visualizer(label_map, encoding="jpg")`, write to
`output_path`).

    python examples/objectdetection/predict.py [--model ssd.pt] [--img_path DIR] [--output_path DIR]

Without `--model` an SSD-300 with fresh weights is used (no pretrained file in this image; train one with ssd_finetune.py), and
without `--img_path` a few pictures of coloured rectangles are written first -- the point is the pipeline."""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import cv2
import numpy as np
import torch

from zoo.feature.common import ChainedPreprocessing
from zoo.feature.image import ImageChannelNormalize, ImageMatToTensor, ImageResize, ImageSet
from zoo.models.image.common.image_config import ImageConfigure
from zoo.models.image.objectdetection import SSDVGG
from zoo.models.image.objectdetection.object_detector import ObjectDetector, ScaleDetection, Visualizer

ap = argparse.ArgumentParser()
ap.add_argument("--model", default=None, help="file written by torch.save(net) / torch.jit.save")
ap.add_argument("--img_path", default=None)
ap.add_argument("--output_path", default=None)
ap.add_argument("--images", type=int, default=3)
args = ap.parse_args()

img_dir = args.img_path or tempfile.mkdtemp(prefix="zoo_od_in_")
out_dir = args.output_path or tempfile.mkdtemp(prefix="zoo_od_out_")
if args.img_path is None:
    rng = np.random.RandomState(0)
    for i in range(args.images):
        mat = rng.randint(0, 60, (240, 320, 3)).astype(np.uint8)
        x1, y1 = rng.randint(0, 200), rng.randint(0, 140)
        mat[y1:y1 + 80, x1:x1 + 100] = [(255, 40, 40), (40, 255, 40), (40, 40, 255)][i % 3]
        cv2.imwrite(os.path.join(img_dir, "img_%02d.png" % i), mat)

label_map = {0: "__background__", 1: "blue box", 2: "green box", 3: "red box"}
config = ImageConfigure(pre_processor=ChainedPreprocessing([ImageResize(300, 300), ImageChannelNormalize(123.0, 117.0, 104.0), ImageMatToTensor()]),
                        post_processor=ScaleDetection(), batch_per_partition=1, label_map=label_map)
if args.model:
    model = ObjectDetector.load_model(args.model)
    model.config = config
else:
    torch.manual_seed(0)
    model = ObjectDetector(SSDVGG(class_num=4, resolution=300, conf_thresh=0.2, keep_topk=20), config=config)

image_set = ImageSet.read(img_dir)
output = model.predict_image_set(image_set)
Visualizer(model.get_config().label_map(), thresh=0.25, encoding="jpg")(output)
n_boxes = 0
for f in output.get_image_features():
    det = np.asarray(f["predict"]).reshape(-1, 6)
    det = det[det[:, 0] >= 0]
    n_boxes += len(det)
    name = os.path.splitext(os.path.basename(f["uri"]))[0] + "_detection.jpg"
    with open(os.path.join(out_dir, name), "wb") as fh:
        fh.write(f["visualized"])
print("detected %d boxes in %d images; visualizations written to %s: %s" % (n_boxes, len(output.get_image_features()), out_dir, sorted(os.listdir(out_dir))))
