"""Decode GEMM (weight streaming, M = 32) in CUDA-graph replay: 24 back-to-back calls on DISTINCT weight matrices (so
nothing is served from L2), per shape and K split (= cluster size): microseconds per call and effective GB/s."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from pretraining_llm_b200 import ops  # noqa: E402

M, REPS = 32, 24
shapes = [("qkv", 6144, 2048), ("mlp up", 8192, 2048), ("mlp down", 2048, 8192)]
for name, N, K in shapes:
    ws = [(torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16) for _ in range(REPS)]
    x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    line = []
    for split in (0, 1, 2, 3, 4, 6, 8):
        try:
            for w in ws[:2]:
                ops.gemm_skinny(x, w, out, split_k=split)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for w in ws:
                    ops.gemm_skinny(x, w, out, split_k=split)
            best = 1e9
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                g.replay()
                e1.record()
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            us = best * 1e3 / REPS
            line.append(f"split {split}: {us:.1f} us {N * K * 2 / us / 1e3:.0f} GB/s")
        except Exception as e:  # noqa: BLE001
            line.append(f"split {split}: {type(e).__name__}")
    print(f"{name} {N}x{K}: " + " | ".join(line), flush=True)
