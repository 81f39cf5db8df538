"""Low-precision image inference through InferenceModel (reference examples: pyzoo/zoo/examples/openvino/predict.py and
pyzoo/zoo/examples/vnni/openvino/predict.py: `InferenceModel.load_tf_object_detection_as_openvino` / `load_openvino(xml, bin)`,
read an image folder with ImageSet, resize + channel order + to-tensor, `model.predict(batch)`, print the top classes).

`load_openvino(xml, bin)` imports the IR graph into torch (`zoo.pipeline.inference.openvino_ir`; no Inference Engine here), so the
model runs like any other: convolutions in bf16 channels_last through cuDNN and, after `quantize(calibration_batch)` -- the
counterpart of the IR's int8 mode --, every Linear / Dense layer in FP8 (E4M3) on the tcgen05 GEMM kernel.  The IR read here is
written first from a torch model with `save_openvino_ir` (no model zoo download on this box)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import cv2
import numpy as np
import torch
import torch.nn as nn

from zoo.feature.common import ChainedPreprocessing
from zoo.feature.image import ImageCenterCrop, ImageChannelNormalize, ImageMatToTensor, ImageResize, ImageSet
from zoo.pipeline.inference import InferenceModel

folder = tempfile.mkdtemp()
rng = np.random.RandomState(0)
for i in range(8):
    cv2.imwrite(os.path.join(folder, "img%d.jpg" % i), rng.randint(0, 255, (240, 320, 3)).astype(np.uint8))

torch.manual_seed(0)
net = nn.Sequential(nn.Conv2d(3, 16, 7, stride=4), nn.ReLU(), nn.Conv2d(16, 32, 3, stride=2), nn.ReLU(), nn.AdaptiveAvgPool2d(2), nn.Flatten(),
                    nn.Linear(128, 256), nn.ReLU(), nn.Linear(256, 1000), nn.Softmax(-1)).eval()
images = ImageSet.read(folder).transform(ChainedPreprocessing([ImageResize(256, 256), ImageCenterCrop(224, 224),
                                                                ImageChannelNormalize(123.0, 117.0, 104.0), ImageMatToTensor()]))
batch = torch.from_numpy(np.stack(images.get_image()))

cuda = torch.cuda.is_available() and not os.environ.get("ZOO_FORCE_CPU")
from zoo.pipeline.inference.openvino_ir import save_openvino_ir  # noqa: E402

xml, weights = save_openvino_ir(nn.Sequential(*[m for m in net if not isinstance(m, nn.AdaptiveAvgPool2d)][:4], nn.AvgPool2d(13, 13), *list(net)[5:]),
                                os.path.join(folder, "model.xml"), (1, 3, 224, 224))
model = InferenceModel(1, "cuda" if cuda else "cpu", dtype=torch.bfloat16 if cuda else None)
model.load_openvino(xml, weights)
print("loaded OpenVINO IR:", os.path.basename(xml), "+", os.path.basename(weights), "(%d KB)" % (os.path.getsize(weights) // 1024))
ref = model.predict(batch.numpy())
model.quantize(batch)                                   # calibration on the batch itself; Linear layers -> FP8
out = model.predict(batch.numpy())
for uri, p in zip([f["uri"] for f in images.get_image_features()], out):
    top = np.argsort(-p)[:3]
    print(os.path.basename(uri), "->", [(int(c), round(float(p[c]), 4)) for c in top])
top5 = lambda p: [set(np.argsort(-r)[:5].tolist()) for r in p]
overlap = np.mean([len(a & b) / 5.0 for a, b in zip(top5(out), top5(ref))])
print("predicted %d images; FP8 vs unquantised: top-5 overlap %.2f, largest probability difference %.4f (fresh weights: scores are nearly flat)" %
      (len(out), overlap, float(np.abs(out - ref).max())))
