import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
M, d = 16384, 768
a = torch.randn(M, d, device="cuda").to(BF16); w = torch.randn(4*d, d, device="cuda").to(BF16)
bias = torch.zeros(4*d, device="cuda"); out = torch.empty(M, 4*d, device="cuda", dtype=BF16)
dy = torch.randn(M, 4*d, device="cuda").to(BF16); g = torch.empty(4*d, d, device="cuda")
dx = torch.empty(M, d, device="cuda", dtype=BF16)
for _ in range(3):
    ops.gemm(a, w, out=out, bias=bias, relu=True)            # fwd  (K-major both)
    ops.gemm(dy, w, out=dx, b_mn=True)                       # dgrad (B MN-major)
    ops.gemm(dy, a, out=g, a_mn=True, b_mn=True)             # wgrad (both MN-major, fp32 out)
torch.cuda.synchronize()
print("ok")
