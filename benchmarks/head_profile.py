"""A few launches of the one-pass classification head (K5b) at the NeuralCF shape (B 65,536, K 128, C 5) for `ncu` captures and
CUDA-event timing.  usage: python benchmarks/head_profile.py [B] [K] [C]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from zoo.ops.head_loss import head_softmax_nll

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
K = int(sys.argv[2]) if len(sys.argv) > 2 else 128
C = int(sys.argv[3]) if len(sys.argv) > 3 else 5
dev = "cuda"
xs = [torch.randn(B, K, device=dev).relu_().to(torch.bfloat16) for _ in range(8)]       # 8 x 16.8 MB: rotates through more than... not L2, see flush
w16 = (torch.randn(C, K, device=dev) * 0.2).to(torch.bfloat16)
bias = torch.zeros(C, device=dev)
labels = torch.randint(1, C + 1, (B,), device=dev)
dw, db = torch.zeros(C, K, device=dev), torch.zeros(C, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for i in range(3):
    head_softmax_nll(xs[i], w16, bias, labels, dw, db, False, None, True, K // 2)
torch.cuda.synchronize()
ts = []
for i in range(10):
    flush.zero_()                                   # L2 flush between timed launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    head_softmax_nll(xs[i % 8], w16, bias, labels, dw, db, False, None, True, K // 2)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
ts.sort()
byts = B * K * 2 * 2 + B * 8
print("head B=%d K=%d C=%d: median %.1f us, best %.1f us (includes one 2-float fill); %.1f MB -> %.0f GB/s at the median" %
      (B, K, C, ts[len(ts) // 2], ts[0], byts / 1e6, byts / ts[len(ts) // 2] / 1e3))
