"""Attention kernel timing without launch gaps: one CUDA graph = 12 "layers" (each with its own qkv / out / dout
buffers, > L2 in total, like the real step), replayed; reports the per-layer time of the forward and of the backward
(delta + key-stationary pass + query-stationary pass)."""
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


# usage: bench_attn.py            -> the two model shapes
#        bench_attn.py hs H B T    -> one custom shape
if len(sys.argv) == 5:
    shapes = (tuple(int(v) for v in sys.argv[1:5]),)
else:
    shapes = ((64, 12, 16, 1024), (128, 16, 8, 2048))
for hs, H, B, T in shapes:
    qkvs = [torch.randn(B * T, 3 * H * hs, device="cuda").to(BF16) for _ in range(L)]
    outs, lses = zip(*[ops.attn_fwd(q, B, T, H, hs) for q in qkvs])
    douts = [torch.randn_like(o) for o in outs]
    dqkv = torch.empty_like(qkvs[0]); delta = torch.empty(B, H, T, device="cuda")
    fl = 4.0 * B * H * T * T * hs / 2
    ms = graph_time(lambda: [ops.attn_fwd(qkvs[i], B, T, H, hs, out=outs[i], lse=lses[i]) for i in range(L)])
    print(f"attn fwd hs{hs} B{B} T{T} H{H}: {ms:.4f} ms/layer  {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
    ms = graph_time(lambda: [ops.attn_bwd(qkvs[i], outs[i], douts[i], lses[i], B, T, H, hs, dqkv=dqkv, delta=delta) for i in range(L)])
    print(f"attn bwd hs{hs} B{B} T{T} H{H}: {ms:.4f} ms/layer  {2.5 * fl / ms / 1e9:.1f} TFLOP/s (5-matmul FLOPs)", flush=True)
