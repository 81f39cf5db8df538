"""K8 attention (forward, and forward + backward) against torch SDPA (its flash / efficient library kernels) at the shapes of the
reference's TransformerLayer / BERT: hidden 768 = 12 heads x 64, L = 77 / 128 / 512.  CUDA events, L2 flushed between iterations.
usage: python benchmarks/attention_bench.py [batch]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from zoo.ops.attention import attention_kernel

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
H, D = 12, 64
flush = torch.empty(160 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e3      # us, median


out = []
for L in (77, 128, 512):
    for causal in (False, True):
        q, k, v = (torch.randn(B, H, L, D, device="cuda", dtype=torch.bfloat16, requires_grad=True) for _ in range(3))
        w = torch.randn(B, H, L, D, device="cuda", dtype=torch.bfloat16)
        ours_f = lambda: attention_kernel(q.detach(), k.detach(), v.detach(), None, causal)
        sdpa_f = lambda: F.scaled_dot_product_attention(q.detach(), k.detach(), v.detach(), is_causal=causal)

        def ours_fb():
            o = attention_kernel(q, k, v, None, causal)
            o.backward(w)
            q.grad = k.grad = v.grad = None

        def sdpa_fb():
            o = F.scaled_dot_product_attention(q, k, v, is_causal=causal)
            o.backward(w)
            q.grad = k.grad = v.grad = None

        r = {"B": B, "H": H, "L": L, "causal": causal, "fwd_us": timeit(ours_f), "sdpa_fwd_us": timeit(sdpa_f), "fwd_bwd_us": timeit(ours_fb),
             "sdpa_fwd_bwd_us": timeit(sdpa_fb)}
        flops_f = 4.0 * B * H * L * L * D * (0.5 if causal else 1.0)
        r["fwd_tflops"] = flops_f / r["fwd_us"] / 1e6
        r["fwd_speedup"] = r["sdpa_fwd_us"] / r["fwd_us"]
        r["fwd_bwd_speedup"] = r["sdpa_fwd_bwd_us"] / r["fwd_bwd_us"]
        out.append(r)
        print(json.dumps(r), flush=True)
