"""CUDA-event timing of the non-GEMM kernels against the library path for the same op (torch / cuDNN), L2 flushed between
iterations.  `--once`: one call of each (for ncu).  Results -> gpurun_out/kernels_bench.json
    K6  persistent LSTM / GRU (fwd + bwd) vs torch.nn.LSTM / GRU (cuDNN)
    K7  fused causal conv (+bias+ReLU, fwd + bwd) vs F.pad + F.conv1d + relu
    K13 serving preprocess vs torch (permute/float/interpolate/normalise)
    K14 top-k(+softmax) vs torch.softmax + torch.topk
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from zoo.ops import lstm_persistent as lp
from zoo.ops.serving_ops import preprocess_images, topk
from zoo.ops.tcn import causal_conv1d

once = "--once" in sys.argv
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
rows = []


def timeit(fn, iters=10):
    fn()
    if once:
        return 0.0
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]


def report(name, shape, ours, lib):
    rows.append({"kernel": name, "shape": shape, "zoo_us": round(ours, 1), "library_us": round(lib, 1), "speedup": round(lib / ours, 2) if ours else None})
    print(rows[-1], flush=True)


for kind, cls in (("lstm", nn.LSTM), ("gru", nn.GRU)):
    for (B, T, Fin, H, L) in [(1024, 50, 8, 32, 2), (256, 24, 8, 64, 1), (1024, 128, 32, 128, 1)]:
        rnn = cls(Fin, H, num_layers=L, batch_first=True).cuda()
        x = torch.randn(B, T, Fin, device="cuda", requires_grad=True)

        def ours():
            out, _ = lp.maybe_run(rnn, x, None)
            out[:, -1].sum().backward()

        def lib():
            out, _ = rnn(x)
            out[:, -1].sum().backward()

        report(f"K6 {kind} fwd+bwd", f"B{B} T{T} F{Fin} H{H} L{L}", timeit(ours), timeit(lib))
        with torch.no_grad():
            report(f"K6 {kind} fwd (inference)", f"B{B} T{T} F{Fin} H{H} L{L}", timeit(lambda: lp.maybe_run(rnn, x, None)), timeit(lambda: rnn(x)))

for (B, C, T, K, d) in [(1024, 30, 64, 3, 4), (256, 64, 256, 7, 16)]:
    x = torch.randn(B, C, T, device="cuda", requires_grad=True)
    w = (torch.randn(C, C, K, device="cuda") * 0.1).requires_grad_(True)
    b = torch.zeros(C, device="cuda", requires_grad=True)

    def ours():
        causal_conv1d(x, w, b, d, True).sum().backward()

    def lib():
        torch.relu(F.conv1d(F.pad(x, ((K - 1) * d, 0)), w, b, dilation=d)).sum().backward()

    report("K7 causal conv+bias+relu fwd+bwd", f"B{B} C{C} T{T} K{K} d{d}", timeit(ours), timeit(lib))

rng = np.random.RandomState(0)
imgs = [rng.randint(0, 256, (256, 256, 3), dtype=np.uint8) for _ in range(64)]
dev_imgs = torch.from_numpy(np.stack(imgs)).cuda()
mean, std = [123.68, 116.779, 103.939], [58.393, 57.12, 57.375]
mt, st = torch.tensor(mean, device="cuda").view(1, 3, 1, 1), torch.tensor(std, device="cuda").view(1, 3, 1, 1)
from zoo.ops.serving_ops import ImageBatcher
import ctypes as C
from zoo.native import lib as nl

bat = ImageBatcher("cuda")
Bn, Cc, _ = bat.stage(imgs)
out = torch.empty(64, 3, 224, 224, dtype=torch.bfloat16, device="cuda")
fa = lambda v: (C.c_float * 4)(*(list(v) + [1.0])[:4])


def ours_pre():
    nl.check(nl.cuda_lib().zoo_serving_preproc(bat.dev.data_ptr(), bat.meta_dev.data_ptr(), bat.meta_dev[Bn:].data_ptr(), out.data_ptr(),
                                               64, 3, 224, 224, 224, 224,
                                               C.cast(fa(mean), C.c_void_p), C.cast(fa(std), C.c_void_p), 1, 1, 1, nl.stream_ptr()), "pre")


def lib_pre():
    x = dev_imgs.permute(0, 3, 1, 2).float()
    x = F.interpolate(x, size=(224, 224), mode="bilinear", align_corners=False)
    return ((x.flip(1) - mt) / st).to(torch.bfloat16)


report("K13 preprocess (kernel only, batch staged)", "64 x 256x256x3 -> 3x224x224 bf16", timeit(ours_pre), timeit(lib_pre))
logits = torch.randn(512, 1000, device="cuda").to(torch.bfloat16)
report("K14 softmax+top5", "512 x 1000 bf16", timeit(lambda: topk(logits, 5, True)), timeit(lambda: torch.softmax(logits.float(), -1).topk(5)))
if not once:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/kernels_bench.json", "w"), indent=1)
