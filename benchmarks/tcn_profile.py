import sys; sys.path.insert(0,'/root/repo')
import torch, torch.nn.functional as F
from torch.profiler import profile, ProfilerActivity
from zoo.ops.tcn import causal_conv1d
B,C,T,K,d = (int(v) for v in (sys.argv[1:6] + ["1024","30","64","3","4"][len(sys.argv)-1:]))
x = torch.randn(B, C, T, device="cuda", requires_grad=True)
w = (torch.randn(C, C, K, device="cuda")*0.1).requires_grad_(True)
b = torch.zeros(C, device="cuda", requires_grad=True)
def ours(): causal_conv1d(x, w, b, d, True).sum().backward()
def lib(): torch.relu(F.conv1d(F.pad(x, ((K-1)*d, 0)), w, b, dilation=d)).sum().backward()
for fn in (ours, lib):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5): fn()
        torch.cuda.synchronize()
    print(fn.__name__)
    for e in sorted(prof.key_averages(), key=lambda e:-e.device_time_total)[:12]:
        print(f"  {e.device_time_total/5:8.1f} us x{e.count/5:.0f} {e.key[:90]}")
