"""A few launches of the whole-tower forward kernel at the NeuralCF shape (65,536 x 256 -> 256 -> 128 -> 64) for `ncu` captures and
CUDA-event timing against the three per-layer launches.  usage: python benchmarks/mlp3_profile.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from zoo.ops.dense import FusedMLP

M = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
mlp = FusedMLP(256, [256, 128, 64], activation="relu").cuda()
xs = [(torch.randn(M, 256, device="cuda") * 0.5).to(torch.bfloat16) for _ in range(6)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for mode in ("1", "0"):
    os.environ["ZOO_FUSED_TOWER"] = mode
    with torch.no_grad():
        for i in range(3):
            mlp(xs[i])
        torch.cuda.synchronize()
        ts = []
        for i in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            mlp(xs[i % 6])
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print("tower forward M=%d, %s: median %.1f us, best %.1f us" % (M, "one kernel" if mode == "1" else "three kernels", ts[5], ts[0]))
