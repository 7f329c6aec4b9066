import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
hs, H, B, T = 64, 12, 16, 1024
qkv = torch.randn(B*T, 3*H*hs, device="cuda").to(BF16)
out, lse = ops.attn_fwd(qkv, B, T, H, hs)
dout = torch.randn_like(out); dqkv = torch.empty_like(qkv); delta = torch.empty(B, H, T, device="cuda")
dq = torch.zeros(B*T, H*hs, device="cuda"); cnt = torch.zeros(B*H*((T+127)//128), device="cuda", dtype=torch.int32)
for _ in range(3):
    ops.attn_bwd(qkv, out, dout, lse, B, T, H, hs, dqkv=dqkv, delta=delta, dq_acc=dq, counters=cnt, impl="tc")
torch.cuda.synchronize()
print("ok")
