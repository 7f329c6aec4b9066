"""LM-head forward GEMM (16384 x 50304 x 768) with / without the fused cross-entropy statistics epilogue."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
M, N, K = 16384, 50304, 768
x = torch.randn(M, K, device="cuda").to(BF16); w = (torch.randn(N, K, device="cuda") * 0.02).to(BF16)
bias = torch.zeros(N, device="cuda"); out = torch.empty(M, N, device="cuda", dtype=BF16)
stats = torch.empty(ops.gemm_stats_entries(N), M, 2, device="cuda")


def t(fn, reps=7):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts)


fl = 2.0 * M * N * K
for name, kw in (("plain", {}), ("bias", dict(bias=bias)), ("bias+stats", dict(bias=bias, stats=stats))):
    ms = t(lambda: ops.gemm(x, w, out=out, **kw))
    print(f"lm head fwd {name}: {ms:.4f} ms {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
