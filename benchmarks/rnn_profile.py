"""A few launches of the persistent LSTM forward at the AutoML shape (B 1024, T 50, H 32, 2 layers) for `ncu` captures."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn

from zoo.ops import lstm_persistent as lp

B, T, H, L = (int(v) for v in (sys.argv[1:5] + ["1024", "50", "32", "2"][len(sys.argv) - 1:]))
rnn = nn.LSTM(8, H, num_layers=L, batch_first=True).cuda()
x = torch.randn(B, T, 8, device="cuda")
with torch.no_grad():
    for _ in range(3):
        lp.maybe_run(rnn, x, None)
torch.cuda.synchronize()
print("ok")
