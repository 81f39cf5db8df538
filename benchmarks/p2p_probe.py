"""Single process, 2 GPUs: NVLink topology, peer copy bandwidth, and the embedding gather reading a table that lives
on the peer (random rows / sequential rows, fp32 and bf16 rows)."""
import ctypes as C
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from zoo.native import lib as nl

print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout[:1500])
a = torch.empty(1 << 28, dtype=torch.float32, device="cuda:0")
b = torch.empty(1 << 28, dtype=torch.float32, device="cuda:1")
for _ in range(2):
    b.copy_(a)
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.set_device(0)
e0.record()
for _ in range(5):
    b.copy_(a)
e1.record()
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
print(f"peer copy 0->1: {5 * a.numel() * 4 / e0.elapsed_time(e1) / 1e6:.1f} GB/s   can_access_peer={torch.cuda.can_device_access_peer(0, 1)}")

lib = nl.cuda_lib()
D, B = 128, 65536 * 4
CASES = ((4_000_000, "fp32", torch.float32, 0), (12_500_000, "fp32", torch.float32, 0), (25_000_000, "fp32", torch.float32, 0),
         (25_000_000, "bf16", torch.bfloat16, 1))
for rows, dt_name, dt, code in CASES:
    local = torch.empty(rows, D, device="cuda:0", dtype=dt).normal_()
    peer = torch.empty(rows, D, device="cuda:1", dtype=dt).normal_()
    out = torch.empty(B, D, dtype=torch.bfloat16, device="cuda:0")
    for where, tab in (("local", local), ("peer", peer)):
        for pattern in ("random", "sequential"):
            ids = (torch.randint(0, rows, (B, 1), device="cuda:0") if pattern == "random" else torch.arange(B, device="cuda:0").reshape(B, 1)).long()
            arr = (C.c_void_p * 1)(tab.data_ptr())
            st = torch.cuda.current_stream().cuda_stream
            def run():
                nl.check(lib.zoo_embedding_gather(ids.data_ptr(), 1, 0, 0, B, 1, 0, arr, rows, None, 0, 1, D, code, out.data_ptr(), D, None, None, st), "g")
            run(); torch.cuda.synchronize()
            e0.record()
            for _ in range(5):
                run()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / 5
            print(f"gather rows={rows/1e6:.1f}M {dt_name} {where:5s} {pattern:10s}: {us:8.1f} us  {B * D * local.element_size() / us / 1e3:7.1f} GB/s")
