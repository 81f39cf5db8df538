"""8-bit inference harness (reference: zoo/.../examples/vnni/bigdl/{Perf,ImageNetEvaluation,Predict}.scala and
examples/vnni/openvino/*: load an int8-quantised model, report throughput / latency over N iterations, top-1 / top-5 on a
validation set, and predictions; published numbers in examples/vnni/bigdl/README.md:106-112).

On B200 the 8-bit tensor-core format is FP8 (E4M3): `InferenceModel.quantize(calibration_batch)` turns every Linear / Dense layer
into an FP8 layer of the hand-written tcgen05 GEMM (per-tensor scales from calibration, fp32 accumulation, bias + activation in
the epilogue, E4M3 hand-over between consecutive layers).  The script measures the bf16 and the FP8 engine on the same model and
batch, and the agreement of their top-1 / top-5 with the fp32 model -- the role VNNI int8 vs fp32 plays in the reference."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.pipeline.inference import InferenceModel

ap = argparse.ArgumentParser()
ap.add_argument("-b", "--batchSize", type=int, default=int(os.environ.get("BATCH", 4096)))
ap.add_argument("-i", "--iteration", type=int, default=int(os.environ.get("ITERS", 200)))
ap.add_argument("--features", type=int, default=1024)
ap.add_argument("--classes", type=int, default=1000)
a = ap.parse_args()
cuda = torch.cuda.is_available() and not os.environ.get("ZOO_FORCE_CPU")
dev = torch.device("cuda" if cuda else "cpu")
if not cuda:
    a.batchSize, a.iteration, a.features, a.classes = min(a.batchSize, 256), min(a.iteration, 5), 128, 40

torch.manual_seed(0)


def make():
    torch.manual_seed(0)
    return nn.Sequential(nn.Linear(a.features, 2048 if cuda else 256), nn.ReLU(), nn.Linear(2048 if cuda else 256, 1024 if cuda else 128), nn.ReLU(),
                         nn.Linear(1024 if cuda else 128, a.classes)).eval()


protos = torch.randn(a.classes, a.features)
labels = torch.randint(0, a.classes, (a.batchSize,))
x = (protos[labels] + 0.35 * torch.randn(a.batchSize, a.features)).to(dev)
fp32 = make().to(dev)
with torch.no_grad():       # give the random classifier something to be right about: a least-squares read-out on the prototypes
    feats = fp32[:-1](protos.to(dev))
    fp32[-1].weight.copy_(torch.linalg.lstsq(feats, torch.eye(a.classes, device=dev)).solution.t())
    fp32[-1].bias.zero_()
    want = fp32(x)


def engine(kind):
    m = InferenceModel(1, dev, dtype=torch.bfloat16 if (cuda and kind != "fp32") else None, use_cuda_graph=cuda)
    net = make()
    net.load_state_dict(fp32.state_dict())
    m.load_torch(net)
    if kind == "fp8":
        m.quantize(x[: min(1024, a.batchSize)])
    return m


def sync():
    if cuda:
        torch.cuda.synchronize()


for kind in ("bf16" if cuda else "fp32", "fp8"):
    m = engine(kind)
    for _ in range(3):
        out = m.predict_tensor(x)
    sync()
    t0 = time.perf_counter()
    for _ in range(a.iteration):
        out = m.predict_tensor(x)
    sync()
    dt = (time.perf_counter() - t0) / a.iteration
    out = out.float()
    top1 = float((out.argmax(1) == want.argmax(1)).float().mean())
    top5 = float((out.topk(5).indices == want.argmax(1, keepdim=True)).any(1).float().mean())
    acc = float((out.argmax(1).cpu() == labels).float().mean())
    print("%-5s batch %d: %.3f ms/batch, %.0f records/s | top-1 agreement with fp32 %.4f, top-5 %.4f | accuracy %.4f"
          % (kind, a.batchSize, dt * 1e3, a.batchSize / dt, top1, top5, acc))
print("fp32 accuracy %.4f" % float((want.argmax(1).cpu() == labels).float().mean()))
