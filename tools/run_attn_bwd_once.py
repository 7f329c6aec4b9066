"""One attention backward (delta + DKV pass + DQ pass) for an ncu capture.  Usage: run_attn_bwd_once.py [hs]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
hs = int(sys.argv[1]) if len(sys.argv) > 1 else 64
H, B, T = (12, 16, 1024) if hs == 64 else (16, 8, 2048)
qkv = torch.randn(B * T, 3 * H * hs, device="cuda").to(BF16)
out, lse = ops.attn_fwd(qkv, B, T, H, hs)
dout = torch.randn_like(out); dqkv = torch.empty_like(qkv); delta = torch.empty(B, H, T, device="cuda")
for _ in range(3):
    ops.attn_fwd(qkv, B, T, H, hs)
    ops.attn_bwd(qkv, out, dout, lse, B, T, H, hs, dqkv=dqkv, delta=delta, impl="tc")
torch.cuda.synchronize()
print("ok")
