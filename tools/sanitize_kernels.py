"""Tiny-shape pass over every kernel family, meant to run under `compute-sanitizer --tool memcheck`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
from pretraining_llm_b200.models import Transformer
BF16 = torch.bfloat16
torch.manual_seed(0)
r = lambda *s: torch.randn(*s, device="cuda").to(BF16)
# GEMMs: all majors / epilogues / kernels, with M, N, K tails
a, b = r(200, 72), r(328, 72)
ops.gemm(a, b, bias=torch.randn(328, device="cuda"), relu=True, two_cta=False)
ops.gemm(a, b, two_cta=True, bn=128)
ops.gemm(r(200, 328), r(328, 72), b_mn=True, two_cta=False, bn=64)
out = torch.zeros(328, 72, device="cuda")
ops.gemm(r(200, 328), r(200, 72), out=out, a_mn=True, b_mn=True, f32_mode=2, split_k=2, two_cta=True, bn=128)
ops.gemm_skinny(r(3, 328), r(200, 328), bias=torch.randn(200, device="cuda"), relu=True, split_k=2)
# attention fwd/bwd (both implementations), T tail
for hs in (64, 128):
    B, T, H = 2, 200, 2
    qkv = r(B * T, 3 * H * hs)
    for impl in ("tc", "mma"):
        o, lse = ops.attn_fwd(qkv, B, T, H, hs, impl=impl)
        ops.attn_bwd(qkv, o, r(B * T, H * hs), lse, B, T, H, hs, impl=impl)
# whole model: train steps + generation (embed, LN, CE, AdamW, decode attention, sampler)
m = Transformer(n_head=4, n_embed=256, context_length=64, vocab_size=1000, N_BLOCKS=2).cuda()
x = torch.randint(0, 1000, (2, 64), device="cuda")
for _ in range(3):
    m.engine.train_step(x, x, 1e-3)
m.generate(x[:, :5], 8)
torch.cuda.synchronize()
print("SANITIZE_RUN_OK")
