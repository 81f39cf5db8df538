"""One forward + backward of the K7 layer at a given shape (for ncu captures): python benchmarks/tcn_once.py B C T K d"""
import sys

sys.path.insert(0, "/root/repo")
import torch

from zoo.ops.tcn import causal_conv1d

B, C, T, K, d = (int(v) for v in (sys.argv[1:6] + ["256", "64", "256", "7", "16"][len(sys.argv) - 1:]))
x = torch.randn(B, C, T, device="cuda", requires_grad=True)
w = (torch.randn(C, C, K, device="cuda") * 0.1).requires_grad_(True)
b = torch.zeros(C, device="cuda", requires_grad=True)
for _ in range(3):
    causal_conv1d(x, w, b, d, True).sum().backward()
torch.cuda.synchronize()
print("ok")
