"""One LM-head forward GEMM with the fused CE-statistics epilogue, for an ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
M, N, K = 16384, 50304, 768
x = torch.randn(M, K, device="cuda").to(BF16); w = (torch.randn(N, K, device="cuda") * 0.02).to(BF16)
bias = torch.zeros(N, device="cuda"); out = torch.empty(M, N, device="cuda", dtype=BF16)
stats = torch.empty(ops.gemm_stats_entries(N), M, 2, device="cuda")
for _ in range(3):
    ops.gemm(x, w, out=out, bias=bias, stats=stats)
torch.cuda.synchronize()
print("ok")
