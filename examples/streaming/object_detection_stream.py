"""Streaming object detection (reference examples: pyzoo/zoo/examples/streaming/objectdetection/streaming_object_detection.py and
zoo/.../examples/streaming/objectdetection/StreamingObjectDetection.scala: a Spark `textFileStream` over a directory, each
micro-batch is a list of image paths -> ImageSet -> `ObjectDetector.predict_image_set` -> `Visualizer` -> files under
`output_path`).  Here the micro-batch loop is a directory watcher with the same batch interval semantics; the producer
(`image_path_writer.py`) runs in a thread so that the script is self-contained."""
import argparse
import os
import sys
import tempfile
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import cv2
import numpy as np
import torch
from image_path_writer import package_path_to_text

from zoo.feature.common import ChainedPreprocessing
from zoo.feature.image import ImageChannelNormalize, ImageMatToTensor, ImageResize, ImageSet
from zoo.feature.image.imageset import ImageFeature
from zoo.models.image.common.image_config import ImageConfigure
from zoo.models.image.objectdetection import SSDVGG
from zoo.models.image.objectdetection.object_detector import ObjectDetector, ScaleDetection, Visualizer

ap = argparse.ArgumentParser()
ap.add_argument("--model", default=None)
ap.add_argument("--img_path", default=None)
ap.add_argument("--streaming_path", default=None)
ap.add_argument("--output_path", default=None)
ap.add_argument("--batch_interval", type=float, default=0.5, help="seconds between looks at the streaming directory")
ap.add_argument("--idle_batches", type=int, default=4, help="stop after this many empty micro-batches in a row")
args = ap.parse_args()

img_dir = args.img_path or tempfile.mkdtemp(prefix="zoo_stream_img_")
stream_dir = args.streaming_path or tempfile.mkdtemp(prefix="zoo_stream_txt_")
out_dir = args.output_path or tempfile.mkdtemp(prefix="zoo_stream_out_")
if args.img_path is None:
    rng = np.random.RandomState(1)
    for i in range(4):
        mat = rng.randint(0, 60, (200, 260, 3)).astype(np.uint8)
        mat[40:120, 30 + 40 * i:110 + 40 * i] = (40, 40, 255)
        cv2.imwrite(os.path.join(img_dir, "frame_%02d.png" % i), mat)

config = ImageConfigure(pre_processor=ChainedPreprocessing([ImageResize(300, 300), ImageChannelNormalize(123.0, 117.0, 104.0), ImageMatToTensor()]),
                        post_processor=ScaleDetection(), batch_per_partition=1, label_map={0: "__background__", 1: "box"})
if args.model:
    model = ObjectDetector.load_model(args.model)
    model.config = config
else:
    torch.manual_seed(0)
    model = ObjectDetector(SSDVGG(class_num=2, resolution=300, conf_thresh=0.3, keep_topk=10), config=config)
visualizer = Visualizer(config.label_map(), thresh=0.3, encoding="jpg")


def read_image_file(path):
    mat = cv2.imread(path)
    return ImageFeature(uri=path, mat=mat, originalSize=mat.shape)


def predict(paths):
    """One micro-batch: the `foreachRDD` body of the reference."""
    output = model.predict_image_set(ImageSet([read_image_file(p) for p in paths]))
    visualizer(output)
    for f in output.get_image_features():
        name = time.strftime("%Y%m%d-%H%M%S-") + os.path.splitext(os.path.basename(f["uri"]))[0] + ".jpg"
        with open(os.path.join(out_dir, name), "wb") as fh:
            fh.write(f["visualized"])


producer = threading.Thread(target=package_path_to_text, args=(stream_dir, img_dir, 2, 0.6), daemon=True)
producer.start()
seen, idle, batches, images = set(), 0, 0, 0
while idle < args.idle_batches:
    time.sleep(args.batch_interval)
    new = sorted(f for f in os.listdir(stream_dir) if f.endswith(".txt") and f not in seen)
    if not new:
        idle += 1
        continue
    idle = 0
    paths = []
    for f in new:
        seen.add(f)
        with open(os.path.join(stream_dir, f)) as fh:
            paths += [line.strip() for line in fh if line.strip()]
    predict(paths)
    batches, images = batches + 1, images + len(paths)
    print("micro-batch %d: %d image paths from %s" % (batches, len(paths), new))
producer.join()
print("stream ended: %d images in %d micro-batches, %d result files in %s" % (images, batches, len(os.listdir(out_dir)), out_dir))
