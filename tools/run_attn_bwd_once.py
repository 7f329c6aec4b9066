"""Attention forward + backward (delta, key-stationary pass, query-stationary pass) for an ncu capture:
GPT-2-small shape (hs = 64, one query tile per CTA) and the 1.3B shape (hs = 128, two query tiles per persistent CTA)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
for hs, H, B, T in ((64, 12, 16, 1024), (128, 16, 8, 2048)):
    qkv = torch.randn(B * T, 3 * H * hs, device="cuda").to(BF16)
    out, lse = ops.attn_fwd(qkv, B, T, H, hs)
    dout = torch.randn_like(out); dqkv = torch.empty_like(qkv); delta = torch.empty(B, H, T, device="cuda")
    for _ in range(2):
        ops.attn_fwd(qkv, B, T, H, hs, out=out, lse=lse)
        ops.attn_bwd(qkv, out, dout, lse, B, T, H, hs, dqkv=dqkv, delta=delta, impl="tc")
torch.cuda.synchronize()
print("ok")
