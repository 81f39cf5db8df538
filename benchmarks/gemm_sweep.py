"""Time of the dense kernels against M (fixed overhead vs asymptotic bandwidth), L2 flushed between iterations.
usage: python benchmarks/gemm_sweep.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from zoo.ops.dense import dense_dgrad, dense_forward, dense_wgrad

flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")


def timeit(fn, n=7):
    fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]


for (K, N) in [(256, 256), (256, 128), (128, 64)]:
    for M in (148 * 128, 2 * 148 * 128, 4 * 148 * 128, 65536, 8 * 148 * 128, 16 * 148 * 128):
        x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(torch.bfloat16)
        b = torch.randn(N, device="cuda")
        dy = torch.randn(M, N, device="cuda").to(torch.bfloat16)
        dw = torch.zeros(N, K, device="cuda")
        y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        dx = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
        xbits = torch.full((M, (K + 31) // 32), 0x55555555, dtype=torch.int32, device="cuda")
        t_f = timeit(lambda: dense_forward(x, w, b, "relu", out=y))
        t_d = timeit(lambda: dense_dgrad(dy, w, out=dx, mask_bits=xbits))
        t_w = timeit(lambda: dense_wgrad(dy, x, dw))
        t_c = timeit(lambda: y.copy_(dy))
        print(f"K={K} N={N} M={M:7d} tiles/CTA={M / 128 / 148:5.2f}  fwd {t_f:6.2f}  dgrad {t_d:6.2f}  wgrad {t_w:6.2f} us   "
              f"fwd GB/s {(M * K + M * N) * 2 / t_f / 1e3:7.1f}   (torch copy of [M,N] bf16: {t_c:6.2f} us)")
