"""A few launches of the K8 attention kernels at the BERT-base shape (B 32, 12 heads, L 512) for `ncu` captures."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from zoo.ops.attention import attention_kernel

B, H, L, D = 32, 12, int(sys.argv[1]) if len(sys.argv) > 1 else 512, 64
q, k, v = (torch.randn(B, H, L, D, device="cuda", dtype=torch.bfloat16, requires_grad=True) for _ in range(3))
w = torch.randn(B, H, L, D, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    o = attention_kernel(q, k, v, None, False)
    o.backward(w)
torch.cuda.synchronize()
print("ok")
