"""Decode attention in CUDA-graph replay: 24 back-to-back calls on DISTINCT KV caches (one per layer, nothing served
from L2), batch 32, 16 heads, head size 128: the one-CTA-per-(row, head) kernel against the split-KV kernel."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from pretraining_llm_b200 import ops  # noqa: E402

B, H, hs, Tmax, REPS = 32, 16, 128, 2048, 24
for L in (64, 272, 1024, 2000):
    kc = [torch.randn(B, H, Tmax, hs, device="cuda").to(torch.bfloat16) for _ in range(REPS)]
    vc = [torch.randn(B, H, Tmax, hs, device="cuda").to(torch.bfloat16) for _ in range(REPS)]
    qkv = torch.randn(B, 3 * H * hs, device="cuda").to(torch.bfloat16)
    out = torch.empty(B, H * hs, device="cuda", dtype=torch.bfloat16)
    pos = torch.tensor([L - 1], device="cuda", dtype=torch.int32)
    scratch = torch.empty(B * H * 8 * (hs + 2), device="cuda")
    tickets = torch.zeros(B * H, device="cuda", dtype=torch.int32)
    res = []
    for name, fn in [("one-CTA", lambda k, v: ops.attn_decode(qkv, k, v, out, B, H, hs, Tmax, pos_dev=pos, append=True))] + \
                    [(f"split {s}", (lambda s: lambda k, v: ops.attn_decode_split(qkv, k, v, out, B, H, hs, Tmax, scratch, tickets, s,
                                                                                 pos_dev=pos, append=True))(s)) for s in (1, 2, 4, 8)]:
        fn(kc[0], vc[0])
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for k, v in zip(kc, vc):
                fn(k, v)
        best = 1e9
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        us = best * 1e3 / REPS
        res.append(f"{name}: {us:.1f} us {B * H * L * hs * 4 / us / 1e3:.0f} GB/s")
    print(f"len {L}: " + " | ".join(res), flush=True)
    del kc, vc
