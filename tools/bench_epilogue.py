"""How much of each GEMM's time is its epilogue?  Same shape with increasingly heavy epilogues (graph of 12 calls)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
L = 12


def graph_time(fn, reps=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fn()
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts) / L


M = 16384
for name, N, K, b_mn in (("qkv fwd", 2304, 768, False), ("mlp up fwd", 3072, 768, False), ("mlp down fwd", 768, 3072, False),
                         ("mlp down dgrad", 3072, 768, True), ("mlp up dgrad", 768, 3072, True)):
    xs = [torch.randn(M, K, device="cuda").to(BF16) for _ in range(L)]
    w = (torch.randn(*((K, N) if b_mn else (N, K)), device="cuda") * 0.02).to(BF16)
    outs = [torch.empty(M, N, device="cuda", dtype=BF16) for _ in range(L)]
    bias = torch.zeros(N, device="cuda")
    bits = torch.zeros(M, (N + 63) // 64, device="cuda", dtype=torch.int64)
    fl = 2.0 * M * N * K
    variants = [("plain", {}), ("bias", dict(bias=bias)), ("bias+relu+bits", dict(bias=bias, relu=True, relu_bits_out=bits)),
                ("relu_bits mask", dict(relu_bits=bits))]
    row = []
    for vn, kw in variants:
        ms = graph_time(lambda: [ops.gemm(xs[i], w, out=outs[i], b_mn=b_mn, **kw) for i in range(L)])
        row.append(f"{vn} {ms * 1e3:.1f} us ({fl / ms / 1e9:.0f} TF)")
    print(f"{name} {M}x{N}x{K}: " + " | ".join(row), flush=True)
