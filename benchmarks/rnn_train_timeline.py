import os, sys
sys.path.insert(0, "/root/repo")
import torch, torch.nn as nn
from torch.profiler import ProfilerActivity, profile
from zoo.ops import lstm_persistent as lp
B, T, H, L = 1024, 50, 32, 2
rnn = nn.LSTM(8, H, num_layers=L, batch_first=True).cuda()
x = torch.randn(B, T, 8, device="cuda", requires_grad=True)
def zoo():
    out, _ = lp.maybe_run(rnn, x, None); out.sum().backward()
def cud():
    out, _ = rnn(x); out.sum().backward()
for name, fn in (("zoo", zoo), ("cudnn", cud)):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn(); torch.cuda.synchronize()
    evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time > 0), key=lambda e: e.time_range.start)
    print(name, len(evs), "kernels, busy", round(sum(e.device_time for e in evs), 1), "us")
    import collections
    c = collections.Counter()
    for e in evs: c[e.name[:70]] += e.device_time
    for k, v in c.most_common(14): print(f"   {v:8.1f}  {k}")
