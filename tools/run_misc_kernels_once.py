"""Every non-GEMM / non-attention kernel at its real shape, once, for an ncu capture:
one training step of a ONE-block model with the GPT-2-small widths (embed, LayerNorm fwd/bwd, CE, column sums, AdamW)
and a few eager decode steps of a one-block model with the 1.3B widths (skinny GEMM, decode attention, sampler, ...)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200.models import Transformer
torch.manual_seed(0)
m = Transformer(n_head=12, n_embed=768, context_length=1024, vocab_size=50304, N_BLOCKS=1).cuda()
x = torch.randint(0, 50304, (16, 1024), device="cuda")
for _ in range(2):
    m.engine.train_step(x, x, 5e-4, use_graph=False)
g = Transformer(n_head=16, n_embed=2048, context_length=2048, vocab_size=50304, N_BLOCKS=1).cuda()
p = torch.randint(0, 50304, (32, 16), device="cuda")
g.engine.generate(p, 4, use_graph=False)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
m.engine.train_step(x, x, 5e-4, use_graph=False)
g.engine.generate(p, 3, use_graph=False)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("ok")
