"""Kernel timeline of one inference pass of the persistent LSTM path vs cuDNN (torch.profiler): where the time of a multi-layer,
small-hidden-size recurrent model goes.  usage: python benchmarks/rnn_timeline.py [B] [T] [H] [L]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn
from torch.profiler import ProfilerActivity, profile

from zoo.ops import lstm_persistent as lp

B, T, H, L = (int(v) for v in (sys.argv[1:5] + ["1024", "50", "32", "2"][len(sys.argv) - 1:]))
rnn = nn.LSTM(8, H, num_layers=L, batch_first=True).cuda()
x = torch.randn(B, T, 8, device="cuda")
for name, fn in (("zoo", lambda: lp.maybe_run(rnn, x, None)), ("cudnn", lambda: rnn(x))):
    with torch.no_grad():
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fn()
            torch.cuda.synchronize()
    evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time > 0), key=lambda e: e.time_range.start)
    t0 = evs[0].time_range.start
    print(f"{name}: {len(evs)} kernels, {evs[-1].time_range.end - t0:.1f} us")
    for e in evs:
        print(f"  + {e.time_range.start - t0:7.1f} us  dur {e.device_time:6.1f}  {e.name[:100]}")
