"""CUDA-event timing of the tcgen05 Dense GEMM family on the NCF tower shapes (L2 flushed between
iterations), with the roofline fraction against MEASURED_PEAKS.json.
usage: python benchmarks/gemm_bench.py [--once]   (--once: a single launch of each, for ncu)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from zoo.ops.dense import dense_dgrad, dense_forward, dense_wgrad

once = "--once" in sys.argv
peaks = {"hbm_gbs": 6538.9, "bf16_tflops": 1670.1}
try:
    peaks.update(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))))
except Exception:
    pass
M = 65536
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
rows = []
for (K, N) in [(256, 256), (256, 128), (128, 64), (128, 5)]:
    x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(torch.bfloat16)
    b = torch.randn(N, device="cuda")
    ldn = (N + 7) // 8 * 8
    dy = torch.randn(M, ldn, device="cuda").to(torch.bfloat16)[:, :N]
    dw = torch.zeros(N, K, device="cuda")
    y = torch.empty(M, ldn, device="cuda", dtype=torch.bfloat16)[:, :N]
    dx = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
    bits = torch.zeros(M, (N + 31) // 32, dtype=torch.int32, device="cuda")
    xbits = torch.full((M, (K + 31) // 32), 0x55555555, dtype=torch.int32, device="cuda")
    ops = {"fwd": (lambda: dense_forward(x, w, b, "relu", out=y), (M * K + M * N + N * K) * 2),
           "fwd_bits": (lambda: dense_forward(x, w, b, "relu", out=y, relu_bits=bits), (M * K + M * N + N * K) * 2 + M * N // 8),
           "dgrad_bits": (lambda: dense_dgrad(dy, w, out=dx, mask_bits=xbits), (M * N + M * K + N * K) * 2 + M * K // 8),
           "dgrad": (lambda: dense_dgrad(dy, w, x, out=dx), (M * N + 2 * M * K + N * K) * 2),
           "wgrad": (lambda: dense_wgrad(dy, x, dw), (M * N + M * K) * 2 + N * K * 4)}
    for name, (fn, nbytes) in ops.items():
        fn()
        if once:
            continue
        ts = []
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        t = sorted(ts)[len(ts) // 2]
        flops = 2.0 * M * N * K
        rows.append({"op": name, "M": M, "K": K, "N": N, "us": round(t, 2), "GBps": round(nbytes / t / 1e3, 1),
                     "hbm_frac_of_measured": round(nbytes / t / 1e3 / peaks["hbm_gbs"], 3),
                     "tflops": round(flops / t / 1e6, 1)})
        print(rows[-1])
torch.cuda.synchronize()
if not once:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/gemm_bench.json", "w"), indent=1)
