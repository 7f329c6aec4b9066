"""A few representative tcgen05 GEMMs for an ncu capture: MLP-up forward (bias + ReLU + sign bits), its dgrad (sign-bit
mask, B MN-major), its wgrad (both MN-major, fp32 out), the LM head with CE statistics, and a square 8192^3."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
M, d, V = 16384, 768, 50304
a = torch.randn(M, d, device="cuda").to(BF16); w = torch.randn(4 * d, d, device="cuda").to(BF16)
bias = torch.zeros(4 * d, device="cuda"); out = torch.empty(M, 4 * d, device="cuda", dtype=BF16)
bits = torch.zeros(M, 4 * d // 64, device="cuda", dtype=torch.int64)
w2 = torch.randn(d, 4 * d, device="cuda").to(BF16); dy = torch.randn(M, d, device="cuda").to(BF16)
dh = torch.empty(M, 4 * d, device="cuda", dtype=BF16); g = torch.empty(4 * d, d, device="cuda")
wv = (torch.randn(V, d, device="cuda") * 0.02).to(BF16); bv = torch.zeros(V, device="cuda")
logits = torch.empty(M, V, device="cuda", dtype=BF16); stats = torch.empty(2 * ((V + 255) // 256), M, 2, device="cuda")
sq = torch.randn(8192, 8192, device="cuda").to(BF16); sqo = torch.empty(8192, 8192, device="cuda", dtype=BF16)
for _ in range(2):
    ops.gemm(a, w, out=out, bias=bias, relu=True, relu_bits_out=bits)   # fwd  (K-major both)
    ops.gemm(dy, w2, out=dh, b_mn=True, relu_bits=bits)                 # dgrad (B MN-major) + ReLU mask
    ops.gemm(dh, a, out=g, a_mn=True, b_mn=True)                        # wgrad (both MN-major, fp32 out)
    ops.gemm(a, wv, out=logits, bias=bv, stats=stats)                   # LM head + CE statistics
    ops.gemm(sq, sq, out=sqo)                                           # square
torch.cuda.synchronize()
print("ok")
